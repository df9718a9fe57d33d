/*
 * oracle/overlap_oracle.c -- T2 oracle ("port"): CPU restatement of ExactOverlapper::overlaps.
 *
 * TEST INFRASTRUCTURE ONLY.  The product path (phasm_b200/) never links, loads or calls
 * this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs do, and only as the checker / reported CPU baseline.
 *
 * What it restates (reference: /root/reference/src/overlapper.cpp):
 *   - entry store: addSequence, overlapper.cpp:22-26 (here: one concatenated byte buffer
 *     + offsets; raw bytes, case-sensitive, exactly like seqan::CharString, overlapper.h:26)
 *   - R1, longest suffix(A)==prefix(B) per ordered pair of distinct entries with
 *     L >= max(min_length,1): the "top of stack" emission, overlapper.cpp:64-91
 *   - R2, every occurrence of a whole entry B (|B| >= min_length) inside another entry:
 *     the containment emission, overlapper.cpp:93-115
 *   - identity by entry index, overlapper.cpp:72-74 and 101-104
 *   - row type (a, b, astart, aend, bstart==0, bend): OverlapT, overlapper.h:17
 *
 * The reference reaches these rows through a generalised enhanced suffix array
 * (overlapper.cpp:33-36) and a suffix-tree DFS (overlapper.cpp:38-147).  That index is 64 B
 * per base and single-threaded, so this restatement derives the SAME multiset through
 * prefix anchors: every row's B starts with B[0:k] (k = min(max(l,1),32)), hence every row
 * is found by looking up A[p:p+k] among the entries' first k bytes and comparing forward.
 *
 * Parity pinning: checked against the compiled reference (oracle/_ref, built from the
 * reference's own sources by oracle/Makefile) in tests/test_oracle.py, and against the
 * golden vectors in tests/golden/ which tests/golden/make_golden.py generated from that
 * compiled reference in the build container.
 *
 * It also counts the data-determined quantities that define the algorithmic bytes of
 * C (canonical candidates), sum of m_c (bases compared up to and
 * including the first mismatch) and sum of 2*ceil(m_c/4).
 *
 * Output rows are sorted by (a, b, astart, aend, bend) on entry INDEX, which is also the
 * deterministic order the CUDA path emits.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

typedef struct {
    uint32_t a, b, astart, aend, bend;
} ovo_row;

typedef struct {
    uint64_t anchors;        /* entries with |b| >= l (and >= 1)                */
    uint64_t positions;      /* A positions with |a|-p >= l                     */
    uint64_t candidates;     /* C of SURVEY 8(d)                                */
    uint64_t compared_bases; /* sum m_c                                         */
    uint64_t operand_bytes;  /* sum 2*ceil(m_c/4)                               */
    uint64_t hits;           /* candidates that matched to the end              */
    uint64_t rows;
} ovo_stats;

typedef struct {
    uint64_t key;
    uint32_t b;
} anchor_t;

static int anchor_cmp(const void *x, const void *y) {
    const anchor_t *p = (const anchor_t *)x, *q = (const anchor_t *)y;
    if (p->key != q->key) return p->key < q->key ? -1 : 1;
    return p->b < q->b ? -1 : (p->b > q->b);
}

static int row_cmp(const void *x, const void *y) {
    const ovo_row *p = (const ovo_row *)x, *q = (const ovo_row *)y;
    if (p->a != q->a) return p->a < q->a ? -1 : 1;
    if (p->b != q->b) return p->b < q->b ? -1 : 1;
    if (p->astart != q->astart) return p->astart < q->astart ? -1 : 1;
    if (p->aend != q->aend) return p->aend < q->aend ? -1 : 1;
    if (p->bend != q->bend) return p->bend < q->bend ? -1 : 1;
    return 0;
}

#define HASH_BASE 0x9E3779B97F4A7C15ull /* odd => polynomial hash mod 2^64 */

static inline uint64_t mix64(uint64_t x) {
    x ^= x >> 31;
    x *= 0x7FB5D329728EA185ull;
    x ^= x >> 27;
    return x;
}

typedef struct {
    uint64_t key;
    uint32_t start, count; /* count == 0 => empty slot */
} slot_t;

typedef struct {
    ovo_row *v;
    size_t n, cap;
} rowvec;

static int rv_push(rowvec *r, ovo_row x) {
    if (r->n == r->cap) {
        size_t nc = r->cap ? r->cap * 2 : 1024;
        ovo_row *nv = (ovo_row *)realloc(r->v, nc * sizeof(ovo_row));
        if (!nv) return -1;
        r->v = nv;
        r->cap = nc;
    }
    r->v[r->n++] = x;
    return 0;
}

typedef struct {
    const uint8_t *seq;
    const uint64_t *off;
    uint32_t M;
    uint64_t l, k, bpow;
    const anchor_t *anc;
    const slot_t *tab;
    uint64_t tsize;
    volatile uint32_t next; /* next block of A entries to claim */
    volatile int fail;
} shared_t;

typedef struct {
    shared_t *sh;
    rowvec *rv;
    uint64_t c_pos, c_cand, c_cmp, c_bytes, c_hits;
} worker_t;

#define A_BLOCK 16u

static void *worker_main(void *arg) {
    worker_t *w = (worker_t *)arg;
    shared_t *sh = w->sh;
    const uint8_t *seq = sh->seq;
    const uint64_t *off = sh->off;
    const uint32_t M = sh->M;
    const uint64_t l = sh->l, k = sh->k, bpow = sh->bpow, tsize = sh->tsize;
    const anchor_t *anc = sh->anc;
    const slot_t *tab = sh->tab;
    rowvec *rv = w->rv;
    /* sp_seen[b] == a+1  <=>  the longest suffix-prefix row for (a,b) is already out */
    uint32_t *sp_seen = (uint32_t *)calloc((size_t)(M ? M : 1), sizeof(uint32_t));
    if (!sp_seen) { sh->fail = 1; return NULL; }
    for (;;) {
        uint32_t a0 = __atomic_fetch_add(&sh->next, A_BLOCK, __ATOMIC_RELAXED);
        if (a0 >= M || sh->fail) break;
        uint32_t a1 = a0 + A_BLOCK < M ? a0 + A_BLOCK : M;
        for (uint32_t a = a0; a < a1; ++a) {
            const uint8_t *A = seq + off[a];
            uint64_t la = off[a + 1] - off[a];
            if (la < l) continue;
            uint64_t npos = la - l + 1; /* positions p with la - p >= l */
            w->c_pos += npos;
            uint64_t h = 0;
            for (uint64_t i = 0; i < k; ++i) h = h * HASH_BASE + (uint64_t)A[i] + 1;
            for (uint64_t p = 0; p < npos; ++p) {
                if (p) h = (h - ((uint64_t)A[p - 1] + 1) * bpow) * HASH_BASE + (uint64_t)A[p + k - 1] + 1;
                uint64_t s = mix64(h) & (tsize - 1);
                while (tab[s].count && tab[s].key != h) s = (s + 1) & (tsize - 1);
                if (!tab[s].count) continue;
                uint64_t rem = la - p;
                for (uint32_t i = 0; i < tab[s].count; ++i) {
                    uint32_t b = anc[tab[s].start + i].b;
                    if (b == a) continue; /* identity by index: overlapper.cpp:72-74 */
                    const uint8_t *B = seq + off[b];
                    uint64_t lb = off[b + 1] - off[b];
                    if (memcmp(A + p, B, k) != 0) continue; /* hash collision: no candidate */
                    ++w->c_cand;
                    uint64_t n = rem < lb ? rem : lb;
                    uint64_t m = k;
                    while (m < n && A[p + m] == B[m]) ++m;
                    uint64_t mc = m < n ? m + 1 : n; /* up to and including first mismatch */
                    w->c_cmp += mc;
                    w->c_bytes += 2 * ((mc + 3) / 4);
                    if (m < n) continue;
                    ++w->c_hits;
                    if (rem >= lb) { /* R2: B wholly inside A, overlapper.cpp:93-115 */
                        ovo_row r = {a, b, (uint32_t)p, (uint32_t)(p + lb), (uint32_t)lb};
                        if (rv_push(rv, r)) sh->fail = 1;
                    }
                    if (rem <= lb && sp_seen[b] != a + 1) { /* R1: smallest p == longest L */
                        sp_seen[b] = a + 1;
                        ovo_row r = {a, b, (uint32_t)p, (uint32_t)la, (uint32_t)rem};
                        if (rv_push(rv, r)) sh->fail = 1;
                    }
                }
            }
        }
    }
    free(sp_seen);
    return NULL;
}

void ovo_free(void *p) { free(p); }

/*
 * seq:  concatenated entry bytes; entry e is seq[off[e] .. off[e+1])
 * Returns 0 on success, -1 on allocation failure.  *rows_out is malloc'd (ovo_free).
 */
int ovo_overlaps(const uint8_t *seq, const uint64_t *off, uint32_t M, uint32_t min_length,
                 int nthreads, ovo_row **rows_out, uint64_t *nrows_out, ovo_stats *st) {
    const uint64_t l = min_length < 1 ? 1 : min_length; /* min_length 0 behaves as 1 */
    const uint64_t k = l < 32 ? l : 32;
    ovo_stats S;
    memset(&S, 0, sizeof S);
    *rows_out = NULL;
    *nrows_out = 0;
    if (nthreads <= 0) {
        long nc = sysconf(_SC_NPROCESSORS_ONLN);
        nthreads = nc > 0 ? (int)nc : 1;
    }
    if ((uint32_t)nthreads > M / A_BLOCK + 1) nthreads = (int)(M / A_BLOCK + 1);

    /* ---- anchors: hash of the first k bytes of every entry long enough to be a B ---- */
    anchor_t *anc = (anchor_t *)malloc((size_t)(M ? M : 1) * sizeof(anchor_t));
    if (!anc) return -1;
    uint64_t na = 0;
    for (uint32_t e = 0; e < M; ++e) {
        uint64_t len = off[e + 1] - off[e];
        if (len < l) continue;
        uint64_t h = 0;
        for (uint64_t i = 0; i < k; ++i) h = h * HASH_BASE + (uint64_t)seq[off[e] + i] + 1;
        anc[na].key = h;
        anc[na].b = e;
        ++na;
    }
    S.anchors = na;
    qsort(anc, na, sizeof(anchor_t), anchor_cmp);

    uint64_t tsize = 16;
    while (tsize < 2 * na + 2) tsize <<= 1;
    slot_t *tab = (slot_t *)calloc(tsize, sizeof(slot_t));
    if (!tab) { free(anc); return -1; }
    for (uint64_t i = 0; i < na;) {
        uint64_t j = i;
        while (j < na && anc[j].key == anc[i].key) ++j;
        uint64_t s = mix64(anc[i].key) & (tsize - 1);
        while (tab[s].count) s = (s + 1) & (tsize - 1);
        tab[s].key = anc[i].key;
        tab[s].start = (uint32_t)i;
        tab[s].count = (uint32_t)(j - i);
        i = j;
    }
    uint64_t bpow = 1; /* HASH_BASE^(k-1) */
    for (uint64_t i = 1; i < k; ++i) bpow *= HASH_BASE;

    rowvec *tv = (rowvec *)calloc((size_t)nthreads, sizeof(rowvec));
    worker_t *wk = (worker_t *)calloc((size_t)nthreads, sizeof(worker_t));
    pthread_t *th = (pthread_t *)calloc((size_t)nthreads, sizeof(pthread_t));
    if (!tv || !wk || !th) { free(anc); free(tab); free(tv); free(wk); free(th); return -1; }
    shared_t sh = {seq, off, M, l, k, bpow, anc, tab, tsize, 0, 0};
    for (int t = 0; t < nthreads; ++t) {
        wk[t].sh = &sh;
        wk[t].rv = &tv[t];
    }
    int started = 0;
    for (int t = 1; t < nthreads; ++t) {
        if (pthread_create(&th[t], NULL, worker_main, &wk[t]) != 0) break;
        ++started;
    }
    worker_main(&wk[0]);
    for (int t = 1; t <= started; ++t) pthread_join(th[t], NULL);
    int fail = sh.fail;
    uint64_t c_pos = 0, c_cand = 0, c_cmp = 0, c_bytes = 0, c_hits = 0;
    for (int t = 0; t < nthreads; ++t) {
        c_pos += wk[t].c_pos; c_cand += wk[t].c_cand; c_cmp += wk[t].c_cmp;
        c_bytes += wk[t].c_bytes; c_hits += wk[t].c_hits;
    }
    free(wk);
    free(th);
    free(anc);
    free(tab);
    if (fail) {
        for (int t = 0; t < nthreads; ++t) free(tv[t].v);
        free(tv);
        return -1;
    }
    size_t total = 0;
    for (int t = 0; t < nthreads; ++t) total += tv[t].n;
    ovo_row *rows = (ovo_row *)malloc((total ? total : 1) * sizeof(ovo_row));
    if (!rows) {
        for (int t = 0; t < nthreads; ++t) free(tv[t].v);
        free(tv);
        return -1;
    }
    size_t w = 0;
    for (int t = 0; t < nthreads; ++t) {
        if (tv[t].n) memcpy(rows + w, tv[t].v, tv[t].n * sizeof(ovo_row));
        w += tv[t].n;
        free(tv[t].v);
    }
    free(tv);
    qsort(rows, total, sizeof(ovo_row), row_cmp);
    S.positions = c_pos;
    S.candidates = c_cand;
    S.compared_bases = c_cmp;
    S.operand_bytes = c_bytes;
    S.hits = c_hits;
    S.rows = total;
    if (st) *st = S;
    *rows_out = rows;
    *nrows_out = total;
    return 0;
}
