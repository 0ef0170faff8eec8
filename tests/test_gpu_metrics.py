"""Row f4: the device-side evaluation summary (deepmetv2_b200/metrics.py, csrc/metrics.cu) against the reference's own
numpy code, evaluate.py:110-156, restated verbatim below as the checker."""
import numpy as np
import pytest
import torch

from deepmetv2_b200 import synth
from deepmetv2_b200.metrics import ResolutionAccumulator
from deepmetv2_b200.model import net as gnet

pytestmark = pytest.mark.gpu
DEV = "cuda"


def reference_summary(resolutions_arr, qT_arr):
    """evaluate.py:110-156 (verbatim arithmetic)."""
    max_x, x_n = 400, 40
    bin_edges = np.arange(0, max_x, 10)
    inds = np.digitize(qT_arr, bin_edges)
    qT_hist = [(bin_edges[i] + bin_edges[i - 1]) / 2. for i in range(1, len(bin_edges))]
    hists = {}
    for key in resolutions_arr:
        R_arr, u_perp_arr, u_par_arr = resolutions_arr[key][2], resolutions_arr[key][0], resolutions_arr[key][1]
        u_perp_hist, u_perp_scaled_hist, u_par_hist, u_par_scaled_hist, R_hist = [], [], [], [], []
        for i in range(1, len(bin_edges)):
            sel = np.where(inds == i)[0]
            R_i = R_arr[sel]
            R_hist.append(np.mean(R_i))
            u_perp_i = u_perp_arr[sel]
            u_perp_scaled_i = u_perp_i / np.mean(R_i)
            u_perp_hist.append((np.quantile(u_perp_i, 0.84) - np.quantile(u_perp_i, 0.16)) / 2.)
            u_perp_scaled_hist.append((np.quantile(u_perp_scaled_i, 0.84) - np.quantile(u_perp_scaled_i, 0.16)) / 2.)
            u_par_i = u_par_arr[sel]
            u_par_scaled_i = u_par_i / np.mean(R_i)
            u_par_hist.append((np.quantile(u_par_i, 0.84) - np.quantile(u_par_i, 0.16)) / 2.)
            u_par_scaled_hist.append((np.quantile(u_par_scaled_i, 0.84) - np.quantile(u_par_scaled_i, 0.16)) / 2.)
        h = lambda w: np.histogram(qT_hist, bins=x_n, range=(0, max_x), weights=w)  # noqa: E731
        hists[key] = {'u_perp_resolution': h(u_perp_hist), 'u_perp_scaled_resolution': h(u_perp_scaled_hist),
                      'u_par_resolution': h(u_par_hist), 'u_par_scaled_resolution': h(u_par_scaled_hist), 'R': h(R_hist)}
    return hists


def test_epoch_summary_matches_evaluate_py():
    torch.manual_seed(0)
    acc = ResolutionAccumulator(DEV, capacity=64)             # small capacity: exercises the growth path
    res_ref, qt_ref = None, np.zeros(0, np.float32)
    g = torch.Generator().manual_seed(5)
    for step in range(6):
        b = synth.make_batch(250, "tiny", seed=100 + step)
        # qT = |y[:, :2]| must populate every bin up to 390 GeV (the reference's np.quantile raises on an empty bin)
        qt_true = torch.rand(250, generator=g) * 392.0
        th = torch.rand(250, generator=g) * 6.2831853
        b.y[:, 0], b.y[:, 1] = qt_true * torch.cos(th), qt_true * torch.sin(th)
        d = b.to(DEV)
        w = torch.rand(d.x.shape[0], generator=g).to(DEV)
        acc.update(w, d.x, d.y, d.batch)
        r, qt = gnet.resolution(w, d.x, d.y, d.batch)         # the per-batch metric (already checked against the fixtures)
        qt_ref = np.concatenate((qt_ref, qt))
        if res_ref is None:
            res_ref = {k: [v[0].copy(), v[1].copy(), v[2].copy()] for k, v in r.items()}
        else:
            for k in res_ref:
                for i in range(3):
                    res_ref[k][i] = np.concatenate((res_ref[k][i], r[k][i]))
    arrays, qt_dev = acc.arrays()
    assert np.array_equal(qt_dev, qt_ref)
    for k in res_ref:
        for i in range(3):
            assert np.array_equal(arrays[k][i], res_ref[k][i])
    with np.errstate(all="ignore"):
        want = reference_summary(res_ref, qt_ref)
    got = acc.summary()
    assert set(got) == set(want)
    inds = np.digitize(qt_ref, np.arange(0, 400, 10))
    assert all((inds == i).sum() > 2 for i in range(1, 40)), "test data must populate every bin"
    for k in want:
        for f in want[k]:
            hw, ew = want[k][f]
            hg, eg = got[k][f]
            assert np.allclose(eg, ew)
            scale = np.abs(hw).max()
            assert np.allclose(hg, hw, rtol=2e-5, atol=2e-6 * scale), (k, f, np.abs(hg - hw).max())


def test_empty_bins_are_nan_like_numpy_mean():
    acc = ResolutionAccumulator(DEV)
    b = synth.make_batch(8, "tiny", seed=3).to(DEV)           # |truth| ~ 30 GeV: high-qT bins stay empty
    acc.update(torch.rand(b.x.shape[0], device=DEV), b.x, b.y, b.batch)
    h = acc.summary()["MET"]["R"][0]
    assert np.isnan(h[35]) and h[39] == 0.0
