"""``python -m myfyp_b200`` -> the CLI (reference: p2pfl/__main__.py:20-23)."""

from myfyp_b200.cli import main

if __name__ == "__main__":
    main()
