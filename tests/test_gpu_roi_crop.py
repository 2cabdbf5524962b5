"""GPU parity of K6 (mouth-ROI crop + gray + resize) -- byte work, so the bar is bit-exact: against crops produced by
the reference's own crop_roi_gray / crop_roi (tests/golden) and against the CPU oracle on seeded boxes."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda")


@pytest.mark.parametrize("name", ["roi_crop.npz", "roi_crop_many.npz"])
def test_roi_crop_vs_reference_outputs(golden_dir, name):
    dev = _cuda()
    from silent_speech_b200 import crop_rois
    g = np.load(os.path.join(golden_dir, name))
    N = len(g["widths"])
    frames = torch.from_numpy(g["frame"])[None].to(dev)
    idx = torch.zeros(N, dtype=torch.int32, device=dev)
    centers = torch.from_numpy(g["centers"]).to(dev)
    for flavor, key, okk, eps in (("live", "roi_live", "ok_live", 0.0), ("record", "roi_record", "ok_record", 1e-6)):
        mw = torch.from_numpy(g["widths"].astype(np.float64) + (eps if name == "roi_crop_many.npz" else 0.0)).to(dev)
        R, ok = crop_rois(frames, centers, mw, flavor=flavor, frame_index=idx)
        assert np.array_equal(ok.cpu().numpy().astype(bool), g[okk]), flavor
        assert np.array_equal(R.cpu().numpy(), g[key]), (flavor, int(np.abs(R.cpu().numpy().astype(int) - g[key]).max()))


@pytest.mark.parametrize("flavor", ["live", "record"])
def test_roi_crop_vs_oracle_seeded(flavor):
    """Camera-sized frames (720x1280), one frame per crop, boxes from the recorder's 60..150 px band plus clipped and
    rejected ones; also checks that frames are addressed per crop."""
    dev = _cuda()
    from oracle import ref_roi_crop as rc
    from silent_speech_b200 import crop_rois
    rng = np.random.default_rng(21)
    F = 12
    frames = rng.integers(0, 256, (F, 720, 1280, 3)).astype(np.uint8)
    frames[:, ::2] //= 2                                     # structure, so that mixing rows/frames up would show
    centers = np.stack([rng.uniform(100, 1180, F), rng.uniform(100, 620, F)], 1).astype(np.float32)
    widths = rng.uniform(60, 150, F)
    centers[0] = [10.0, 5.0]; centers[1] = [1279.0, 719.5]; widths[2] = 80.0; centers[2] = [640.0, 360.0]   # clipped, clipped, 192x160
    widths[3] = 0.4                                          # rejected by live (x2 <= x1+2), tiny for record
    widths = np.array([float(np.float32(w)) for w in widths])
    R, ok = crop_rois(torch.from_numpy(frames).to(dev), torch.from_numpy(centers).to(dev),
                      torch.from_numpy(widths).to(dev), flavor=flavor)
    R, ok = R.cpu().numpy(), ok.cpu().numpy()
    for i in range(F):
        o = rc.crop_roi(flavor, frames[i], centers[i], float(widths[i]))
        assert bool(ok[i]) == (o is not None), i
        want = o if o is not None else np.zeros((48, 96), np.uint8)
        assert np.array_equal(R[i], want), (i, int(np.abs(R[i].astype(int) - want).max()))
