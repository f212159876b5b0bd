#!/usr/bin/env python
"""Does this box expose NVSwitch multicast (NVLS) to the process?  Prints the report of b2_multicast_probe (runtime.cu).

    python tools/nvls_probe.py [> profiles/nvls_probe.txt]
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from myfyp_b200.ops import native  # noqa: E402

lib = native.load()
buf = ctypes.create_string_buffer(8192)
rc = lib.b2_multicast_probe(buf, 8192)
print(buf.value.decode(), end="")
print(f"result: {'multicast objects can be created' if rc == 1 else 'NVLS multicast is not usable from this process' if rc == 0 else 'driver error'} (rc={rc})")
