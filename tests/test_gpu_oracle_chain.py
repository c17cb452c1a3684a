"""The checker chain on the GPU box: the reference KATs against both checkers and the C restatement against the
compiled reference (oracle/_ref travels with the snapshot).  Same test bodies as tests/test_oracle_kats.py and
tests/test_oracle_vs_ref.py (which run in the CPU suite); re-exported here under the gpu marker so that the
driver's `-m gpu` run carries the whole chain restatement -> reference -> CUDA path."""
import pytest

import test_oracle_kats as _kats
import test_oracle_vs_ref as _diff

pytestmark = pytest.mark.gpu

for _mod in (_kats, _diff):
    for _name in dir(_mod):
        if _name.startswith("test_"):
            _fn = getattr(_mod, _name)
            if getattr(_mod, "pytestmark", None) is not None and _mod is _diff:
                _fn = _mod.pytestmark(_fn) if callable(_mod.pytestmark) else _fn
            globals()[f"test_chain_{_mod.__name__.split('_', 1)[1]}_{_name[5:]}"] = _fn
