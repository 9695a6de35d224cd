import sys

from determined_b200.cli.cli import main

sys.exit(main())
