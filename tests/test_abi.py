"""CPU-side checks of the drop-in boundary: libgraphmet.so loads, exports exactly the symbols
include/graphmet.h declares, and the Python host fails loudly without a CUDA device."""
import ctypes
import os
import re

import pytest
import torch

from deepmetv2_b200 import _lib


def _header_symbols():
    src = open(_lib.HEADER).read()
    return sorted(set(re.findall(r"\b(gm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    if not os.path.exists(_lib.SO_PATH):
        _lib.build()
    l = ctypes.CDLL(_lib.SO_PATH)
    names = _header_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(l, n), f"{n} declared in include/graphmet.h but not exported"
    assert sorted(_lib.SIGNATURES) == names, "ctypes SIGNATURES out of sync with the header"
    assert _lib.lib().gm_version() == 100
    assert _lib.lib().gm_status_string(-2) == b"caller-provided buffer too small"


def test_struct_layout_matches_header():
    # 23 trainable tensor pointers + BN buffers: GmParams is 9 + 5 + 4 + 4 + 4*5 + 4 pointers + 2 ints
    assert ctypes.sizeof(_lib.GmBatchNorm) == 5 * 8
    assert ctypes.sizeof(_lib.GmParams) == (9 + 5 + 4 + 4 + 20 + 4) * 8 + 8


def test_state_dict_and_parameter_order_match_reference():
    from deepmetv2_b200.model import net as gnet
    from oracle import graphmet_oracle as O
    ours, ref = gnet.Net(8, 3), O.Net(8, 3)
    assert list(ours.state_dict().keys()) == list(ref.state_dict().keys())
    assert [tuple(p.shape) for p in ours.parameters()] == [tuple(p.shape) for p in ref.parameters()]
    ours.load_state_dict(ref.state_dict(), strict=True)
    opt = torch.optim.AdamW(ours.parameters(), lr=1e-3)
    assert len(opt.param_groups[0]["params"]) == 23


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only check")
def test_no_cpu_fallback():
    from deepmetv2_b200 import ops
    from deepmetv2_b200.model import net as gnet
    with pytest.raises(RuntimeError):
        ops.radius_graph(torch.zeros(4, 2), 0.4)
    with pytest.raises(RuntimeError):
        gnet.loss_fn(torch.zeros(4), torch.zeros(4, 11), torch.zeros(1, 11), torch.zeros(4, dtype=torch.long))
    m = gnet.Net(8, 3)
    with pytest.raises(RuntimeError):
        m(torch.zeros(4, 8), torch.zeros(4, 3, dtype=torch.long), torch.zeros(2, 0, dtype=torch.long),
          torch.zeros(4, dtype=torch.long))
