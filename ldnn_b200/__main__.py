"""`python -m ldnn_b200 --topology ring --balance disbalanced --synthetic …` — same as a variant's main.py."""
import sys

from .cli import main

if __name__ == "__main__":
    main(sys.argv[1:])
