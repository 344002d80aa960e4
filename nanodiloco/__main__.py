"""Entry point of `python -m nanodiloco` (the reference's nanodiloco/__main__.py:1-4): runs the drop-in
training loop of nanodiloco_b200 with the reference's command-line flags."""
import sys

import nanodiloco_b200.main as _loop

if __name__ == "__main__":
    sys.exit(_loop.main())
