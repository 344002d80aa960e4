"""`python -m nanodiloco` (reference nanodiloco/__main__.py:1-4)."""
from nanodiloco_b200.main import main

if __name__ == "__main__":
    main()
