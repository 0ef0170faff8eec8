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


def test_ctypes_structs_match_the_c_header_field_by_field(tmp_path):
    """Every ctypes.Structure of the binding against the C compiler's view of include/graphmet.h: sizeof and the
    offset of every field (a drifted field order or type would hand the kernels garbage pointers)."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no C compiler")
    structs = {n: c for n, c in vars(_lib).items()
               if isinstance(c, type) and issubclass(c, ctypes.Structure) and n.startswith("Gm")}
    assert {"GmBatchNorm", "GmParams", "GmAdamW", "GmDrnConv", "GmDrnParams", "GmDrnView"} <= set(structs)
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{_lib.HEADER}"', 'int main(void) {']
    for n, c in structs.items():
        lines.append(f'  printf("{n} %zu\\n", sizeof({n}));')
        for f in c._fields_:
            lines.append(f'  printf("{n}.{f[0]} %zu\\n", offsetof({n}, {f[0]}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c99", "-o", str(exe), str(src)], check=True)
    want = dict(l.split() for l in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for n, c in structs.items():
        assert ctypes.sizeof(c) == int(want[n]), n
        for f in c._fields_:
            assert getattr(c, f[0]).offset == int(want[f"{n}.{f[0]}"]), f"{n}.{f[0]}"


def test_ctypes_signatures_match_the_header_prototypes():
    """Argument count and scalar / pointer kind of every entry point: _lib.SIGNATURES against the prototypes of
    include/graphmet.h (an argument dropped on one side shifts every pointer after it)."""
    src = re.sub(r"/\*.*?\*/", " ", open(_lib.HEADER).read(), flags=re.S)
    protos = dict((m.group(2), (m.group(1).strip(), m.group(3)))
                  for m in re.finditer(r"^([A-Za-z_][\w \*]*?)\b(gm_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", src, flags=re.M))

    def kind(decl):
        decl = decl.strip()
        if "*" in decl:
            return "ptr"
        base = re.sub(r"\b(const|unsigned|signed)\b", "", decl).split()
        t = " ".join(base[:-1]) if len(base) > 1 else base[0]
        return {"int": "int", "float": "float", "double": "double", "size_t": "size_t", "long long": "ll",
                "int64_t": "ll"}[t]

    ckind = {ctypes.c_int: "int", ctypes.c_float: "float", ctypes.c_double: "double", ctypes.c_size_t: "size_t",
             ctypes.c_longlong: "ll", ctypes.c_void_p: "ptr", ctypes.c_char_p: "ptr"}
    assert sorted(protos) == sorted(_lib.SIGNATURES)
    for name, (ret, params) in protos.items():
        params = params.strip()
        want = [] if params in ("", "void") else [kind(p) for p in params.split(",")]
        got = [ckind.get(a, "ptr") for a in _lib.SIGNATURES[name][1]]      # POINTER(struct) -> ptr
        assert got == want, f"{name}: binding {got} vs header {want}"
        rk = "ptr" if "*" in ret else {"int": "int", "size_t": "size_t", "long long": "ll", "float": "float",
                                       "double": "double", "void": "void"}[re.sub(r"\bconst\b", "", ret).strip()]
        res = _lib.SIGNATURES[name][0]
        assert ("void" if res is None else ckind.get(res, "ptr")) == rk, f"{name}: return type"
