"""Reference arm of bench.py: the UNMODIFIED reference (p2pfl 0.4.0) from ``baseline/_ref``.

Placeholder until the offline install is wired: prints the `unavailable` line and exits 0.
"""

from __future__ import annotations

import json


def run_reference(args) -> None:
    print(json.dumps({"impl": "reference", "unavailable": "reference arm not wired yet"}), flush=True)
