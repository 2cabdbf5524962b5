"""Sliding-window stream scheduling (inactive/live_feed.py:155-213 policy).  CPU: the host integer state against a
literal deque simulation.  GPU: every prediction of the scheduler against the oracle run on the window the
reference loop would have built."""
import collections

import numpy as np
import pytest
import torch


def _simulate(S, W, pred_every, warmup, ticks, rng):
    """The reference loop, literally: per stream a deque(maxlen=W) of frame tags."""
    bufs = [collections.deque(maxlen=W) for _ in range(S)]
    frame_idx = np.zeros(S, int)
    for t in range(ticks):
        ids = np.sort(rng.choice(S, size=int(rng.integers(1, S + 1)), replace=False))
        found = rng.uniform(size=ids.size) > 0.2
        events = []
        for i, f in zip(ids, found):
            frame_idx[i] += 1
            if not f:
                continue                                    # live_feed.py:178-184: "No face found" -> next camera frame
            bufs[i].append((t, int(i)))
            if len(bufs[i]) >= warmup and frame_idx[i] % pred_every == 0:
                events.append((int(i), list(bufs[i])))
        yield t, ids, found, events


def test_window_state_matches_deque_simulation():
    from silent_speech_b200.streaming import WindowState
    S, W = 7, 12
    st = WindowState(S, W, pred_every=2, warmup_min=5)
    ring = {}                                              # (stream, pos) -> tag: what ssb_ring_push would store
    rng = np.random.default_rng(0)
    n_events = 0
    for t, ids, found, events in _simulate(S, W, 2, 5, 120, rng):
        a, pos, d, start, cnt = st.advance(ids.astype(np.int64), found)
        for i, p in zip(a, pos):
            ring[(int(i), int(p))] = (t, int(i))
        assert [int(x) for x in d] == [e[0] for e in events]
        for i, s0, c, (_, want) in zip(d, start, cnt, events):
            got = [ring[(int(i), int((s0 + k) % W))] for k in range(int(c))]
            assert got == want                              # oldest -> newest, exactly the deque's content
            n_events += 1
    assert n_events > 100
    with pytest.raises(ValueError):
        st.advance(np.array([1, 1]), np.array([True, True]))


@pytest.mark.gpu
@pytest.mark.parametrize("use_roi", [True, False])
def test_stream_scheduler_vs_oracle(use_roi):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    dev = torch.device("cuda")
    from oracle import ref_model
    from silent_speech_b200 import BiGRUClassifier, synth
    from silent_speech_b200.streaming import StreamScheduler
    S, W, ticks = 6, 16, 44
    p = ref_model.synth_params(31, use_roi=use_roi)
    m = BiGRUClassifier(180, 10, use_roi=use_roi, precision="tc3")
    m.load_state_dict(p)
    m = m.to(dev).eval()
    sch = StreamScheduler(m, S, max_t=W, pred_every=2, warmup_min=5)
    rng = np.random.default_rng(3)
    d = synth.clip_batch(S, ticks, mode="full", seed=8, use_roi=use_roi)       # per-stream frame sequences
    bufs = [collections.deque(maxlen=W) for _ in range(S)]
    frame_idx = np.zeros(S, int)
    checked = 0
    for t in range(ticks):
        ids = np.sort(rng.choice(S, size=int(rng.integers(2, S + 1)), replace=False))
        found = rng.uniform(size=ids.size) > 0.15
        X = torch.from_numpy(d["X"][ids, t]).to(dev)
        R = torch.from_numpy(d["R"][ids, t]).to(dev) if use_roi else None
        due, logits = sch.tick(ids, X, R, face_found=found)
        want_ids = []
        for i, f in zip(ids, found):
            frame_idx[i] += 1
            if not f:
                continue                                    # no face: no append, no prediction (live_feed.py:178-184)
            bufs[i].append(t)
            if len(bufs[i]) >= 5 and frame_idx[i] % 2 == 0:
                want_ids.append(int(i))
        assert [int(x) for x in due] == want_ids
        if not want_ids:
            assert logits is None
            continue
        got = logits.cpu()
        for row, i in enumerate(want_ids):
            ts = list(bufs[i])
            Xw = np.zeros((1, W, 180), np.float32); Xw[0, :len(ts)] = d["X"][i, ts]
            Rw = None
            if use_roi:
                Rw = np.zeros((1, W, 48, 96), np.uint8); Rw[0, :len(ts)] = d["R"][i, ts]
            with torch.no_grad():
                ref = ref_model.forward(p, torch.from_numpy(Xw), torch.tensor([len(ts)]), None if Rw is None else torch.from_numpy(Rw))
            err = float((got[row] - ref[0]).abs().max() / ref.abs().max())
            assert err < 1e-4, (t, i, err)
            assert int(got[row].argmax()) == int(ref[0].argmax())
            checked += 1
    assert checked > 30
