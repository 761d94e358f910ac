"""python -m raytracing_b200: the non-interactive mirror of RT_bench.py's menus (cli.py)."""
import sys

from .cli import main

sys.exit(main())
