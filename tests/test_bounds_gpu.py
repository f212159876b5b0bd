"""Guard-band run of every kernel family (tools/bounds_check.py): NaN-poisoned guards around inputs, canaries around outputs.

compute-sanitizer is closed on the B200 pool (profiles/sanitizer/README.md); this is the bounds evidence that can be produced there."""
import importlib.util
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not torch.cuda.is_available(), reason="needs a GPU")
def test_no_kernel_reads_or_writes_outside_its_buffers():
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "bounds_check.py")
    spec = importlib.util.spec_from_file_location("bounds_check", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.run(verbose=False)
