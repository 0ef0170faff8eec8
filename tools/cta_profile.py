"""Per-CTA busy cycles of the geo kernels (needs a libgraphmet built with -DGEO_PROFILE; GRAPHMET_SO=...).
    python tools/cta_profile.py [c2|c5]"""
import ctypes as C, os, sys
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from deepmetv2_b200 import synth, _lib
from deepmetv2_b200.model import net as gnet
from deepmetv2_b200.trainer import GraphMETTrainer

wl = sys.argv[1] if len(sys.argv) > 1 else "c5"
n_ev, shape = {"c2": (32, "dytt"), "c3": (256, "highpu"), "c5": (32, "highpu")}[wl]
torch.manual_seed(0)
model = gnet.Net(8, 3).to("cuda")
tr = GraphMETTrainer(model, lr=1e-3)
b = synth.make_batch(n_ev, shape, seed=1000 * int(wl[1])).to("cuda")
men = int((b.ptr[1:] - b.ptr[:-1]).max())
for _ in range(3):
    tr.train_step(b, men)
torch.cuda.synchronize()
l = _lib.lib()
buf = np.zeros((4, 2048), dtype=np.int64)

fn = C.CDLL(_lib.SO_PATH).gm_debug_cta_cycles
fn.argtypes = [C.c_void_p]
assert fn(buf.ctypes.data) == 0
for k, name in ((0, "k_geo_aggr"), (1, "k_geo_scatter")):
    v = buf[k]; v = v[v > 0]
    print(f"{name}: CTAs {len(v)}  mean {v.mean():.0f}  max {v.max()}  min {v.min()}  max/mean {v.max() / v.mean():.3f}  "
          f"p90/mean {np.percentile(v, 90) / v.mean():.3f}")
