import sys

from phasm.cli.assembler import main

sys.exit(main())
