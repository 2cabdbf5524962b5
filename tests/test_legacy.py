"""the earlier model generation (GRUClassifier + the shipped trained weights word_model_5.pt) and the
mouth-open EMA / hysteresis detector, against the fixture written by oracle/make_golden.py::golden_legacy (reference
classes run with the reference's checkpoint) and against torch's nn.GRU for other hidden sizes."""
import os

import numpy as np
import pytest
import torch

from silent_speech_b200 import legacy

GOLD = os.path.join(os.path.dirname(__file__), "golden", "legacy.npz")


def _state_dict(g):
    return {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd_")}


def test_state_dict_names_and_pad_or_trim():
    g = np.load(GOLD)
    sd = _state_dict(g)
    m = legacy.GRUClassifier(83, 5, hidden=64)
    assert set(m.state_dict().keys()) == set(sd.keys())
    m.load_state_dict(sd)                                            # shapes of the shipped checkpoint
    X = np.arange(12, dtype=np.float32).reshape(4, 3)
    assert legacy.pad_or_trim(X, 6).shape == (6, 3) and (legacy.pad_or_trim(X, 6)[4:] == 0).all()
    assert (legacy.pad_or_trim(X, 2) == X[:2]).all()
    with pytest.raises(ValueError):
        legacy.GRUClassifier(83, 5, hidden=192)
    with pytest.raises(Exception):
        m(torch.zeros(1, 60, 83))                                    # CPU tensors: no CPU path


@pytest.mark.gpu
def test_word_model_5_matches_reference():
    g = np.load(GOLD)
    m = legacy.GRUClassifier(83, 5, hidden=64)
    m.load_state_dict(_state_dict(g))
    m = m.cuda().eval()
    x = torch.from_numpy(g["x"]).cuda()
    logits = m(x)
    torch.cuda.synchronize()
    np.testing.assert_allclose(logits.cpu().numpy(), g["logits"], rtol=0, atol=2e-6)
    assert (logits.argmax(1).cpu().numpy() == g["logits"].argmax(1)).all()
    # the recurrence output and the fused time mean themselves
    from silent_speech_b200 import ops
    H, gr = 64, m.gru
    gi = torch.empty(x.shape[0], x.shape[1], 6 * H, device="cuda")
    x2 = x.view(-1, 83)
    ops.linear_nt(x2, gr.weight_ih_l0, gr.bias_ih_l0, gi.view(-1, 6 * H), c_col0=0)
    ops.linear_nt(x2, gr.weight_ih_l0_reverse, gr.bias_ih_l0_reverse, gi.view(-1, 6 * H), c_col0=3 * H)
    pooled, out = ops.gru_small_fwd(gi, gr.weight_hh_l0, gr.bias_hh_l0, gr.weight_hh_l0_reverse, gr.bias_hh_l0_reverse,
                                    hidden=H, want_out=True)
    np.testing.assert_allclose(out.cpu().numpy(), g["out"], rtol=0, atol=2e-6)
    np.testing.assert_allclose(pooled.cpu().numpy(), g["pooled"], rtol=0, atol=2e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("H,D,B,T,bidir,packed", [(32, 20, 3, 9, True, False), (96, 50, 37, 25, True, True),
                                                  (128, 33, 65, 12, True, False), (64, 83, 40, 60, False, True),
                                                  (64, 7, 1, 1, True, True)])
def test_gru_small_vs_torch(H, D, B, T, bidir, packed):
    """Generic shapes against torch's nn.GRU (the arithmetic the reference class calls), un-packed and packed."""
    from silent_speech_b200 import ops
    torch.manual_seed(H + D)
    ref = torch.nn.GRU(D, H, 1, batch_first=True, bidirectional=bidir)
    x = torch.randn(B, T, D)
    lens = torch.randint(1, T + 1, (B,))
    lens[0] = T
    with torch.no_grad():
        if packed:
            pk = torch.nn.utils.rnn.pack_padded_sequence(x, lens, batch_first=True, enforce_sorted=False)
            o, _ = ref(pk)
            want, _ = torch.nn.utils.rnn.pad_packed_sequence(o, batch_first=True, total_length=T)
        else:
            want, _ = ref(x)
    nd = 2 if bidir else 1
    gi = torch.empty(B, T, nd * 3 * H, device="cuda")
    xd = x.cuda().view(-1, D)
    p = {k: v.detach().cuda() for k, v in ref.named_parameters()}
    ops.linear_nt(xd, p["weight_ih_l0"], p["bias_ih_l0"], gi.view(-1, nd * 3 * H), c_col0=0)
    if bidir:
        ops.linear_nt(xd, p["weight_ih_l0_reverse"], p["bias_ih_l0_reverse"], gi.view(-1, nd * 3 * H), c_col0=3 * H)
    pooled, out = ops.gru_small_fwd(gi, p["weight_hh_l0"], p["bias_hh_l0"], p.get("weight_hh_l0_reverse"),
                                    p.get("bias_hh_l0_reverse"), lengths=lens.cuda() if packed else None, hidden=H, want_out=True)
    torch.cuda.synchronize()
    np.testing.assert_allclose(out.cpu().numpy(), want.numpy(), rtol=0, atol=3e-6)
    np.testing.assert_allclose(pooled.cpu().numpy(), want.mean(dim=1).numpy(), rtol=0, atol=3e-6)


@pytest.mark.gpu
def test_mouth_open_detector_matches_reference_loop():
    g = np.load(GOLD)
    S, F = g["m_valid"].shape
    lm = np.zeros((S, F, 478, 3), np.float32)
    lm[:, :, g["m_ids"], :2] = g["m_lm4"]
    alpha, open_thr, close_thr = (float(v) for v in g["m_const"])
    det = legacy.MouthOpenDetector(S, alpha=alpha, open_thr=open_thr, close_thr=close_thr)
    lmd, vd = torch.from_numpy(lm).cuda(), torch.from_numpy(g["m_valid"]).cuda()
    # two calls: the state carries across calls exactly like the reference's loop variables
    e1, o1 = det.update(lmd[:, :33].contiguous(), vd[:, :33].contiguous())
    e2, o2 = det.update(lmd[:, 33:].contiguous(), vd[:, 33:].contiguous())
    ema = torch.cat([e1, e2], 1).cpu().numpy()
    opn = torch.cat([o1, o2], 1).cpu().numpy()
    np.testing.assert_allclose(ema, g["m_ema"], rtol=1e-13, atol=0)
    np.testing.assert_array_equal(opn, g["m_open"])
    np.testing.assert_allclose(det.ema.cpu().numpy(), g["m_ema"][:, -1], rtol=1e-13)
