"""Self-play generation on the GPU (config C5, one rank): games are played by the device kernels (choices, tensor-core
evaluation, policy sampling, update), recorded and packed on the device. Checks, on the produced .battle.data bytes:
  * every game replays on the ORACLE from its stored initial battle with its stored choices to exactly the stored
    result and length — the reference's own consistency check (cpp/include/py/battle/encoded-frames.h:117-126);
  * the stored choices are legal and the stored policies/values are the network's (oracle evaluation of the replayed
    state, u16-quantised as compress_probs does);
  * files are flushed by size, finished slots are refilled with fresh games."""
import json
import os

import numpy as np
import pytest

from oak_ref import ROOT, NetOracle, Oracle, make_net_blob, parse_battle_data

pytestmark = pytest.mark.gpu


def test_selfplay_frames_replay_on_oracle(tmp_path):
    from oak_b200.selfplay import FrameWriter, SelfPlay
    orc = Oracle()
    pool = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
    blob = make_net_blob(np.random.default_rng(77), scale=3.0)
    sp = SelfPlay(768, blob, pool, seed=1234, max_updates=400)
    w = FrameWriter(str(tmp_path), buffer_mb=1)
    done = sp.run(260, writer=w, collect_every=4)
    w.flush()
    sp.close()
    assert done > 768 and sp.started >= 768 + done  # every finished slot was refilled
    assert len(w.files) >= 2
    stream = b"".join(open(f, "rb").read() for f in w.files)
    games = parse_battle_data(stream)
    assert len(games) == done
    net = NetOracle(orc, blob)
    checked = 0
    for gi, g in enumerate(games):
        b, d = g["battle"].copy(), np.zeros(8, np.uint8)
        r = 0x50  # both players move after the leads are out (PKMN::update(0, 0) result)
        assert len(g["updates"]) >= 1
        for ui, u in enumerate(g["updates"]):
            ch1 = orc.choices(b, 0, (r >> 4) & 3)
            ch2 = orc.choices(b, 1, (r >> 6) & 3)
            assert (u["m"], u["n"]) == (len(ch1), len(ch2)) and u["c1"] in ch1 and u["c2"] in ch2 and u["iterations"] == 0
            if gi % 25 == 0 and ui % 7 == 0:
                v, l1, l2, _ = net.eval_one(b, d, ch1, ch2)
                for lg, got in ((l1, u["p1_emp"]), (l2, u["p2_emp"])):
                    p = np.exp(lg.astype(np.float32))
                    p = p / p.sum(dtype=np.float32)
                    assert np.abs((p.astype(np.float64) * 65535.0).astype(np.int64) - got.astype(np.int64)).max() <= 2
                assert abs(int(v * 65535.0) - u["ev"]) <= 2 and u["ev"] == u["nv"] and np.array_equal(u["p1_emp"], u["p1_nash"])
                checked += 1
            r, _ = orc.update(b, u["c1"], u["c2"], d)
            assert (r & 15) == 0 or ui + 1 == len(g["updates"])
        assert r == g["result"] and (r & 15) != 0
    assert checked > 50


def test_device_drain_and_gather_equal_the_host_drain():
    """SelfPlay.collect(device=True) + sharding.gather_streams (the path the 8-GPU generator takes: the drained stream stays
    in device memory until the writer rank has it) delivers the bytes the host drain delivers: two generators with the
    same seed, one drained each way."""
    import torch
    from oak_b200.selfplay import SelfPlay
    from oak_b200.sharding import gather_streams
    pool = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
    blob = make_net_blob(np.random.default_rng(5), scale=3.0)
    a = SelfPlay(512, blob, pool, seed=42, max_updates=400)
    b = SelfPlay(512, blob, pool, seed=42, max_updates=400)
    from oak_b200.sharding import StreamGatherer
    gatherer, previous = StreamGatherer(device=torch.device("cuda:0")), None
    total = 0
    for s in range(160):
        a.step()
        b.step()
        if (s + 1) % 8 == 0:
            sa, ga = a.collect()
            sb, gb = b.collect(device=True)
            assert torch.is_tensor(sb) and sb.is_cuda and ga == gb
            got = gather_streams(sb)  # one rank: the writer's own copy
            assert len(got) == 1 and np.array_equal(got[0], sa)
            late = gatherer.submit(sb)  # the asynchronous form hands back the PREVIOUS drain
            if previous is not None:
                assert late is not None and len(late) == 1 and np.array_equal(late[0], previous)
            previous = sa
            total += ga
    last = gatherer.flush()
    assert np.array_equal(last[0], previous) and gatherer.flush() is None
    assert total > 100
    a.close()
    b.close()


def test_device_drains_reach_the_writer_through_the_page_locked_ring():
    """The collective-free frame path (sharding.FrameSender / FrameCollector): device drains are copied by DMA into this
    rank's page-locked shared-memory ring and the writer thread reads them — the same bytes, in order, as the host drain of
    a twin generator; a small ring (wraps + back-pressure) and the drain buffer being reused at once are both exercised."""
    import torch
    from oak_b200.selfplay import SelfPlay
    from oak_b200.sharding import FrameCollector, FrameSender
    pool = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
    blob = make_net_blob(np.random.default_rng(5), scale=3.0)
    a = SelfPlay(512, blob, pool, seed=43, max_updates=400)
    b = SelfPlay(512, blob, pool, seed=43, max_updates=400)
    tag = f"gputest{os.getpid()}"
    sender = FrameSender(0, device=torch.device("cuda:0"), capacity=1 << 20, tag=tag)
    assert sender.ring.pinned is not None and sender.ring.pinned.is_pinned()
    arrived = []
    collector = FrameCollector(1, lambda r, view: arrived.append(view.tobytes()), tag=tag)
    want = []
    for s in range(240):
        a.step()
        b.step()
        if (s + 1) % 8 == 0:
            sa, ga = a.collect()
            sb, gb = b.collect(device=True)
            assert ga == gb
            if len(sa):
                want.append(bytes(sa))
            sender.submit(sb)
            sb.fill_(0xEE)  # the caller's buffer is free again as soon as submit returns
    sender.flush()
    collector.wait_idle()
    collector.stop()
    assert arrived == want and sum(map(len, want)) > (1 << 20)  # more bytes than the ring holds: it wrapped
    assert collector.bytes == sender.bytes
    sender.close()
    assert not [f for f in os.listdir("/dev/shm") if tag in f]
    a.close()
    b.close()


def test_pipelined_lanes_generate_what_separate_generators_do():
    """PipelinedSelfPlay: two lanes on their own streams, one lane's drain + refill under the other's queued steps. Each
    lane's drained bytes are exactly those of a lone SelfPlay with the lane's seed and game ids, and replay on the oracle."""
    import torch
    from oak_b200.selfplay import PipelinedSelfPlay, SelfPlay
    pool = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
    blob = make_net_blob(np.random.default_rng(5), scale=3.0)
    psp = PipelinedSelfPlay(384, blob, pool, seed=77, lanes=2, first_game=1000, max_updates=400)
    got = {0: [], 1: []}

    def on_drain(stream, games):
        lane = [l for l, st in enumerate(psp.streams) if st == torch.cuda.current_stream()][0]
        got[lane].append(stream.cpu().numpy().copy())
    done = psp.run(200, on_drain, collect_every=8)
    torch.cuda.synchronize()
    assert done > 100
    for lane in (0, 1):
        lone = SelfPlay(384, blob, pool, seed=77 + 7919 * lane, first_game=1000 + (lane << 28), max_updates=400)
        want = []
        for s in range(200):
            lone.step()
            if (s + 1) % 8 == 0:
                want.append(np.array(lone.collect()[0], copy=True))
        lone.close()
        assert np.array_equal(np.concatenate(got[lane]), np.concatenate(want)), lane
    games = parse_battle_data(np.concatenate(got[0] + got[1]).tobytes())
    assert len(games) == done
    orc = Oracle()
    for g in games[::7]:  # whole games that replay to their stored result
        b, d, r = g["battle"].copy(), np.zeros(8, np.uint8), 0x50
        for u in g["updates"]:
            r, _ = orc.update(b, u["c1"], u["c2"], d)
        assert r == g["result"] and (r & 15) != 0
    psp.close()
