#!/usr/bin/env python3
"""Benchmark of the hot path: random playouts of lockstep Generation I battles (BASELINE.json configs[1]).

One "step" = one pass of the hot path over one batch: 2^20 synthetic sample-team battles per GPU
 played to a terminal result by the fused rollout kernel (choices x2 + pick + update per
battle-step). `value` = battle-steps/s of the whole job with inputs resident in HBM; `e2e` = the same
through the C ABI with host buffers (upload of battle/durations/result/seed records and download of
score/turns/steps inside the timed region).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--battles B] [--impl reference]
  torchrun --nproc-per-node N bench.py --gpus N ...      (one rank per GPU)

--impl reference times the reference's CPU algorithm (the oracle restatement of libpkmn + Oak's playout
loop; the Zig toolchain is unavailable so oracle/_ref cannot be built) on all host cores.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

B_ROLLOUT = 404  # algorithmic bytes per fused rollout: read 384+8+1+8, write 1+2
B_STEP = 787     # algorithmic bytes per battle-step for the one-launch-per-step kernel
F_EVAL = 1048384      # dense-equivalent FLOP per leaf evaluation
F_EVAL_TC = 2625536   # FLOP per evaluation the tensor-core kernel actually issues: 32 * sum(N * MMAs) over its chunk
                      # program, i.e. three products per split-bf16 GEMM and the K/N padding included


def issue_stats():
    """Issue-slot / lane-efficiency figures of the rollout kernel from the newest ncu summary tools/ncu_issue.py wrote under
    profiles/ (with its capture date and commit), or None: nothing is quoted that was not measured."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_rollout_issue.json")))
    if not files:
        return None
    doc = json.load(open(files[-1]))
    doc["file"] = os.path.relpath(files[-1], ROOT)
    doc["note"] = "from a separate ncu capture (see captured / commit / command), not re-measured in this run"
    return doc


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return json.load(open(path)), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0]))
                mx.append(float(c[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_rollouts(orc, first, n, threads):
    from oak_ref import oracle_rollout_batch, oracle_sample_batch, sample_seeds
    b, d, r = oracle_sample_batch(orc, first, n, threads)
    seeds = sample_seeds(first, n)
    t0 = time.perf_counter()
    _, _, steps = oracle_rollout_batch(orc, b, d, r, seeds, threads=threads)
    dt = time.perf_counter() - t0
    return int(steps.astype(np.int64).sum()), dt


def run_reference(args, rank, world):
    """The reference arm: its CPU playout loop (oracle restatement) on all host cores."""
    if rank != 0:
        return
    from oak_ref import Oracle
    orc = Oracle()
    threads = os.cpu_count() or 1
    n = args.ref_battles
    for w in range(args.warmup):
        cpu_rollouts(orc, 0, min(n, 1 << 14), threads)
    total_steps, total_t, total_roll = 0, 0.0, 0
    for k in range(args.steps):
        s, dt = cpu_rollouts(orc, k * n, n, threads)
        total_steps += s
        total_t += dt
        total_roll += n
    v = total_steps / total_t
    line = {
        "impl": "reference", "metric": "battle-steps/sec", "value": v, "unit": "battle-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "rollouts_per_s": total_roll / total_t,
        "config": {"workload": "random playouts to terminal of synthetic RBY OU sample-team battles (BASELINE configs[1])",
                   "battles_per_gpu": n, "parallelism": "all host cores, one playout stream per thread",
                   "note": "the GPU arm's step: 2^20 playouts of the same synthetic battles (successive steps take successive battle indices)"},
        "cpu_baseline": {"value": v, "unit": "battle-steps/s", "cores": threads, "kind": "port",
                         "sample": f"{args.steps} x {n} playouts, CPU restatement of libpkmn/Oak (Zig toolchain unavailable)"},
        "e2e": {"value": v, "unit": "battle-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else a library prints (NCCL's version banner, ...) was
    sent to stderr by main()."""
    os.write(_STDOUT_FD, (json.dumps(line) + "\n").encode())


_STDOUT_FD = 1


def main():
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)  # native libraries writing to fd 1 now land on stderr
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--battles", type=int, default=1 << 20, help="battles per GPU per step")
    ap.add_argument("--impl", default="oak_b200", choices=["oak_b200", "reference"])
    ap.add_argument("--ref-battles", type=int, default=1 << 20, help="playouts per step of the reference arm (the GPU arm's step)")
    ap.add_argument("--cpu-sample", type=int, default=1 << 23)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import oak_b200
    from oak_ref import sample_seeds, sample_teams

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; oak_b200 has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    n = args.battles
    from oak_b200.sharding import shard_range
    first, _ = shard_range(rank, world, n)  # rank r owns synthetic battles [r*n, (r+1)*n): no data-path collective
    eng = oak_b200.BattleEngine(n, device=local)
    stream = torch.cuda.current_stream(dev)
    eng.set_stream(stream.cuda_stream)

    # ---- inputs: built on the device (PKMN::battle + update(0,0)), then kept resident
    seeds_np = sample_seeds(first, n)
    eng.build_battles(sample_teams(first, n), seeds_np)
    pb, pd, pr, _ = eng.resident()
    hb, hd, hr = eng.download()  # host copies for the e2e leg (pinned below)
    seeds = torch.from_numpy(seeds_np.view(np.int64)).to(dev)
    score = torch.empty(n, dtype=torch.uint8, device=dev)
    turns = torch.empty(n, dtype=torch.int16, device=dev)
    steps_t = torch.empty(n, dtype=torch.int32, device=dev)

    class Raw:  # device pointer wrapper for the resident arrays
        def __init__(self, p, n):
            self.p, self.shape = p, (n,)

        def data_ptr(self):
            return self.p
    rb, rd, rr = Raw(pb, n), Raw(pd, n), Raw(pr, n)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # The K steps are issued alternately on two CUDA streams (two contexts over the same resident arrays, separate outputs):
    # a launch ends with a ~5 ms tail in which only its longest playouts are still running (profiles/r1p_rollout_tail.txt), and
    # the next launch's CTAs take the SMs it frees. Every step is a full launch over all n battles; the timed region is
    # first launch -> both streams drained.
    stream_b = torch.cuda.Stream(dev)
    eng_b = oak_b200.BattleEngine(1, device=local)
    eng_b.set_stream(stream_b.cuda_stream)
    outs = [(score, turns, steps_t), (torch.empty_like(score), torch.empty_like(turns), torch.empty_like(steps_t))]
    lanes = [(eng, stream), (eng_b, stream_b)]

    def one_step(k=0):
        e_, _ = lanes[k & 1]
        e_.rollout_device(rb, rd, rr, seeds, 100000, 0, *outs[k & 1])

    for k in range(args.warmup):
        one_step(k)
    barrier()
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record(stream)
    one_step(0)  # one launch with the GPU to itself (untimed warm-up step): its duration includes the serial tail
    a1.record(stream)
    barrier()
    alone_ms = a0.elapsed_time(a1)
    launches0 = eng.launch_count + eng_b.launch_count
    sampler = ClockSampler(local) if rank == 0 else None
    ev_start, ev_end, ev_b = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    ev_start.record(stream)
    stream_b.wait_event(ev_start)
    for k in range(args.steps):
        st_ = lanes[k & 1][1]
        evs[k][0].record(st_)
        one_step(k)
        evs[k][1].record(st_)
    ev_b.record(stream_b)
    stream.wait_event(ev_b)
    ev_end.record(stream)
    barrier()
    clocks = sampler.stop() if sampler else None
    launches = eng.launch_count + eng_b.launch_count - launches0
    ms_total = ev_start.elapsed_time(ev_end)
    kernel_ms = [a.elapsed_time(b) for a, b in evs]  # per launch, on the launch's own stream (launches overlap at their ends)
    steps_per_pass = int(steps_t.to(torch.int64).sum().item())
    assert args.steps < 2 or torch.equal(outs[0][2], outs[1][2]), "the two streams' launches disagree"
    eng_b.close()

    # ---- e2e through the C ABI with pinned HOST buffers (upload + rollout + download per step)
    phb = torch.from_numpy(hb).pin_memory()
    phd = torch.from_numpy(hd).pin_memory()
    phr = torch.from_numpy(hr).pin_memory()
    phs = torch.from_numpy(seeds_np.view(np.int64)).pin_memory()
    # Two host threads, each with its own context and output buffers, take the steps alternately (the C calls release the
    # GIL): one call's upload and serial tail run under the other call's playouts, as the two streams of the resident leg do.
    import threading
    from oak_b200.engine import BattleEngine as _BE
    e2e_engs = [eng, _BE(n, device=local)]
    e2e_out = [(torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.int16).pin_memory(),
                torch.empty(n, dtype=torch.int32).pin_memory()) for _ in e2e_engs]
    lib, h = eng.lib, eng._h

    def e2e_step(k):
        e_, (o_score, o_turns, o_steps) = e2e_engs[k], e2e_out[k]
        rc = lib.oakb200_rollout_host(e_._h, n, phb.data_ptr(), phd.data_ptr(), phr.data_ptr(), phs.data_ptr(), 100000,
                                      o_score.data_ptr(), o_turns.data_ptr(), o_steps.data_ptr())
        if rc:
            raise RuntimeError(lib.oakb200_last_error(e_._h).decode())

    def e2e_run(total):
        errs = []

        def worker(k, count):
            try:
                for _ in range(count):
                    e2e_step(k)
            except Exception as ex:  # surfaced below
                errs.append(ex)
        ts_ = [threading.Thread(target=worker, args=(k, (total + 1 - k) // 2)) for k in range(2)]
        for t_ in ts_:
            t_.start()
        for t_ in ts_:
            t_.join()
        if errs:
            raise errs[0]

    e2e_iters = max(2, args.steps)
    e2e_run(2)
    barrier()
    l0 = sum(e_.launch_count for e_ in e2e_engs)
    t0 = time.perf_counter()
    e2e_run(e2e_iters)
    barrier()
    e2e_s = time.perf_counter() - t0
    launches += sum(e_.launch_count for e_ in e2e_engs) - l0
    assert torch.equal(e2e_out[0][2], e2e_out[1][2]) and torch.equal(e2e_out[0][0], e2e_out[1][0]), "the two e2e contexts disagree"
    e2e_steps = int(e2e_out[0][2].to(torch.int64).sum().item())
    e2e_engs[1].close()

    # ---- secondary: the one-launch-per-battle-step kernel (lockstep API), HBM-resident, 16 steps
    lock_b = torch.from_numpy(hb).to(dev)
    lock_d = torch.from_numpy(hd).to(dev)
    lock_r = torch.from_numpy(hr).to(dev)
    for s in range(3):
        eng.step_random_device(lock_b, lock_d, lock_r, seeds, s)
    torch.cuda.synchronize(dev)
    le0, le1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    live_before = int(((lock_r & 15) == 0).sum().item())
    le0.record(stream)
    LOCK = 16
    for s in range(3, 3 + LOCK):
        eng.step_random_device(lock_b, lock_d, lock_r, seeds, s)
    le1.record(stream)
    torch.cuda.synchronize(dev)
    lock_ms = le0.elapsed_time(le1) / LOCK
    launches += LOCK + 3

    # ---- secondary: batched leaf evaluation (BASELINE configs[2]: batch 64K, default network dims)
    from oak_ref import make_net_blob
    EV = min(1 << 16, n)
    from oak_b200.sharding import broadcast_bytes, gather_streams
    # rank 0 owns the parameter image; every other rank receives it over NCCL, then loads it
    blob = broadcast_bytes(make_net_blob(np.random.default_rng(1234)).tobytes() if rank == 0 else b"", src=0, device=dev)
    eng.load_network(blob)
    ev_v = torch.empty(EV, dtype=torch.float32, device=dev)
    ev_l = torch.empty((EV, 2, 9), dtype=torch.float32, device=dev)
    ev_c = torch.empty((EV, 2), dtype=torch.uint8, device=dev)
    eb, ed, er = lock_b[:EV], lock_d[:EV], lock_r[:EV]  # states after 19 random-playout steps
    eval_ms_mode = {}
    EVIT = 10
    for mode in (1, 2):  # 1 = fp32 CUDA-core kernel, 2 = tcgen05 tensor-core kernel (the default path)
        eng.set_eval_mode(mode)
        for _ in range(3):
            eng.evaluate_device(eb, ed, er, ev_v, ev_l, ev_c)
        torch.cuda.synchronize(dev)
        ee0, ee1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ee0.record(stream)
        for _ in range(EVIT):
            eng.evaluate_device(eb, ed, er, ev_v, ev_l, ev_c)
        ee1.record(stream)
        torch.cuda.synchronize(dev)
        eval_ms_mode[mode] = ee0.elapsed_time(ee1) / EVIT
        launches += EVIT + 3
    eng.set_eval_mode(0)
    eval_ms_one = eval_ms_mode[2]
    # The same batches issued alternately on two CUDA streams (a second context with the same parameters, its own outputs),
    # as the two self-play lanes and the search's root groups issue them: 65 536 evaluations are 512 tiles on 148 SMs, i.e.
    # 3.46 rounds run as 4, and the next launch's CTAs take the 80 SMs that have no fourth tile.
    ev_eng_b = _BE(1, device=local)
    ev_stream_b = torch.cuda.Stream(dev)
    ev_eng_b.set_stream(ev_stream_b.cuda_stream)
    ev_eng_b.load_network(blob)
    ev_out_b = (torch.empty_like(ev_v), torch.empty_like(ev_l), torch.empty_like(ev_c))
    ev_lanes = [(eng, stream, (ev_v, ev_l, ev_c)), (ev_eng_b, ev_stream_b, ev_out_b)]
    EV2 = 20
    for k in range(4):
        ev_lanes[k & 1][0].evaluate_device(eb, ed, er, *ev_lanes[k & 1][2])
    torch.cuda.synchronize(dev)
    es0, es1, esb = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    es0.record(stream)
    ev_stream_b.wait_event(es0)
    for k in range(EV2):
        ev_lanes[k & 1][0].evaluate_device(eb, ed, er, *ev_lanes[k & 1][2])
    esb.record(ev_stream_b)
    stream.wait_event(esb)
    es1.record(stream)
    torch.cuda.synchronize(dev)
    eval_ms = es0.elapsed_time(es1) / EV2
    launches += EV2 + 4
    assert torch.equal(ev_v, ev_out_b[0]) and torch.equal(ev_c, ev_out_b[2]), "the two evaluation streams disagree"
    ev_eng_b.close()
    # end to end through the C ABI with pinned HOST buffers: records in, value/logits/counts out
    peb, ped, per_ = (x.cpu().pin_memory() for x in (eb, ed, er))
    pev = torch.empty(EV, dtype=torch.float32).pin_memory()
    pel = torch.empty((EV, 2, 9), dtype=torch.float32).pin_memory()
    pec = torch.empty((EV, 2), dtype=torch.uint8).pin_memory()

    def eval_e2e(k=0):
        e_, (o_v, o_l, o_c) = ev_engs[k], ev_out[k]
        rc = lib.oakb200_eval_host(e_._h, EV, peb.data_ptr(), ped.data_ptr(), per_.data_ptr(), o_v.data_ptr(), o_l.data_ptr(), o_c.data_ptr())
        if rc:
            raise RuntimeError(lib.oakb200_last_error(e_._h).decode())
    # one caller: the calls follow one another, each ends with its own download
    ev_engs, ev_out = [eng], [(pev, pel, pec)]
    EV_ONE = 50
    for _ in range(3):
        eval_e2e()
    t0 = time.perf_counter()
    for _ in range(EV_ONE):
        eval_e2e()
    eval_e2e_one_s = (time.perf_counter() - t0) / EV_ONE
    launches += EV_ONE + 3
    # two callers (host threads with a context each, as two search threads of one process would be): one call's upload
    # runs under the other's kernels and download, so the PCIe link stays busy. This is the leg's headline figure.
    ev2 = _BE(EV, device=local)
    ev2.load_network(blob)
    ev_engs.append(ev2)
    ev_out.append((torch.empty(EV, dtype=torch.float32).pin_memory(), torch.empty((EV, 2, 9), dtype=torch.float32).pin_memory(),
                   torch.empty((EV, 2), dtype=torch.uint8).pin_memory()))

    def eval_e2e_run(total):
        errs = []

        def worker(k, count):
            try:
                for _ in range(count):
                    eval_e2e(k)
            except Exception as ex:  # surfaced below
                errs.append(ex)
        ts_ = [threading.Thread(target=worker, args=(k, total // 2)) for k in range(2)]
        for t_ in ts_:
            t_.start()
        for t_ in ts_:
            t_.join()
        if errs:
            raise errs[0]
    EV_CALLS = 200  # ~0.1 s: long enough that thread start-up does not show
    eval_e2e_run(10)
    l0 = sum(e_.launch_count for e_ in ev_engs)
    t0 = time.perf_counter()
    eval_e2e_run(EV_CALLS)
    eval_e2e_s = (time.perf_counter() - t0) / EV_CALLS
    launches += sum(e_.launch_count for e_ in ev_engs) - l0
    assert torch.equal(ev_out[0][0], ev_out[1][0]) and torch.equal(ev_out[0][2], ev_out[1][2]), "the two evaluation contexts disagree"
    ev2.close()
    # ---- secondary: self-play generation (BASELINE configs[4]): lockstep games, network policy, frames recorded and
    # packed on the device, drained per rank into page-locked shared-memory rings the writer (rank 0) reads — no collective
    import json as _json
    from oak_b200.selfplay import SelfPlay
    pool = np.array(_json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
    # two lanes of 65 536 games on their own streams, software-pipelined: one lane's drain + refill (host work that waits
    # for that lane's kernels) runs under the other lane's queued steps (oak_b200.selfplay.PipelinedSelfPlay)
    from oak_b200.selfplay import PipelinedSelfPlay
    SP_LANES, SP_LANE_G, SP_STEPS = 2, 65536, 96
    SP_G = SP_LANES * SP_LANE_G
    psp = PipelinedSelfPlay(SP_LANE_G, blob, pool, seed=99 + rank, device=local, lanes=SP_LANES, first_game=rank << 32, max_updates=512)
    sp = psp.lanes[0]

    # frames reach the writer without a collective (per-GPU form): every rank DMAs its drains over its own PCIe
    # link into a page-locked shared-memory ring (oak_b200.sharding.FrameSender); rank 0's writer thread reads all rings
    from oak_b200.sharding import FrameCollector, FrameSender
    sender = FrameSender(rank, device=dev, capacity=256 << 20)
    barrier()  # start-up only: every ring exists before the writer attaches
    arrived, keeping, at_writer = [], [False], [0] * world  # writer rank: bytes per rank; the streams in the untimed check pass

    def writer_sink(r, view):
        at_writer[r] += len(view)
        if keeping[0]:
            arrived.append(np.array(view, copy=True))
    collector = FrameCollector(world, writer_sink) if rank == 0 else None

    sp_phases = [0.0] * 4

    def sp_pass(keep_arrivals=False):
        # every drain goes to this rank's ring as it is produced; the copy runs behind the generator's next steps
        streams, games = [], 0
        arrived.clear()
        keeping[0] = keep_arrivals
        for r_ in range(world):
            at_writer[r_] = 0
        def on_drain(st_, g_):  # the drained stream stays on the device; the DMA takes it to the ring
            streams.append(st_.clone())
            sender.submit(st_)
        tp = [time.perf_counter()]
        games = psp.run(SP_STEPS, on_drain, collect_every=8)
        torch.cuda.synchronize(dev)
        tp.append(time.perf_counter())
        sender.flush()
        tp.append(time.perf_counter())
        barrier()  # every rank has published its last drain
        tp.append(time.perf_counter())
        if collector is not None:
            collector.wait_idle()  # ... and the writer has consumed it: the timed region ends with the bytes at the writer
        tp.append(time.perf_counter())
        sp_phases[:] = [round((tp[k + 1] - tp[k]) * 1e3, 2) for k in range(4)]  # generate, flush, barrier, writer (ms, this rank)
        return streams, games, sum(at_writer)

    sp_pass()  # warm-up: the same pass
    barrier()
    l0 = psp.launch_count
    sp_runs = []
    for _ in range(3):  # three timed passes, the median is reported (the host side of this leg varies from pass to pass)
        barrier()
        t0 = time.perf_counter()
        sp_streams, sp_games, sp_bytes = sp_pass()
        barrier()
        sp_runs.append((time.perf_counter() - t0, list(sp_phases)))
    sp_runs.sort(key=lambda x: x[0])
    sp_s, sp_phases_timed = sp_runs[1]
    sp_all_s = [round(x[0] * 1e3, 2) for x in sp_runs]
    local_stream = torch.cat(sp_streams).cpu().numpy() if sp_streams else np.empty(0, np.uint8)
    launches += psp.launch_count - l0
    # outside the timed region: the bytes that ARRIVED at the writer (all ranks' drains, read from their rings) are whole
    # games that replay on the device to their stored results, and are as many bytes as the ranks drained
    chk_streams, _, chk_bytes = sp_pass(keep_arrivals=True)  # one more pass, untimed, whose arrivals are kept and replayed
    sent_t = torch.tensor([float(sender.bytes)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(sent_t, op=dist.ReduceOp.SUM)  # after the timed region: how many bytes the ranks sent in total
    sent_total = float(sent_t.item())
    barrier()
    gather_ok = None
    if rank == 0:
        from oak_b200.training_rows import BattleFrames
        collector.stop()
        got_bytes = sum(len(x) for x in arrived)
        gather_ok = got_bytes == chk_bytes and got_bytes > 0 and collector.bytes == int(sent_total)
        if gather_ok:
            # every arrived message is whole games: replay them in slices the engine's capacity (n battles) holds
            from oak_b200.engine import rows_index_host as _rih
            bf, at = None, 0
            try:
                while at < len(arrived) and gather_ok:
                    part, part_rows = [], 0
                    while at < len(arrived):
                        _r = _rih(arrived[at])[2]
                        if part and part_rows + _r > n // 2:
                            break
                        part.append(arrived[at])
                        part_rows += _r
                        at += 1
                    if bf is None or bf.size < part_rows:
                        bf = BattleFrames(max(part_rows, n // 2), device=local, engine=eng)
                    gather_ok = bf.fill(np.concatenate(part)) == part_rows  # raises if a game does not replay to its result
            except ValueError:
                gather_ok = False
            del bf
        if world == 1:
            gather_ok = gather_ok and got_bytes == sum(int(x.numel()) for x in chk_streams)
    barrier()
    sender.close()

    # ---- secondary: training rows (pyoak EncodedBattleFrames.from_bytes / sample): the games this rank's
    # self-play finishes are replayed on the device and encoded into the learner's dense tensors. Input: the self-play
    # stream itself (more untimed steps until ~2^21 stored updates are there), so the whole chain is product kernels.
    from oak_b200.engine import rows_index_host
    from oak_b200.training_rows import EncodedBattleFrames
    TR_ROWS = 1 << 21
    tr_parts, tr_have = [local_stream], 0
    with torch.cuda.stream(psp.streams[0]):
        for s_ in range(400):
            if tr_have >= TR_ROWS:
                break
            sp.step()
            if (s_ + 1) % 8 == 0:
                st_, g_ = sp.collect()
                tr_parts.append(np.array(st_, copy=True))
                tr_have = rows_index_host(np.concatenate(tr_parts))[2] if g_ else tr_have
        psp.streams[0].synchronize()
    psp.close()
    tr_stream = np.concatenate(tr_parts)
    tr_off, tr_base, tr_rows = rows_index_host(tr_stream)
    if tr_rows > TR_ROWS:  # whole games only
        cut = int(np.searchsorted(tr_base, TR_ROWS, side="right")) - 1
        tr_stream, tr_off, tr_rows = tr_stream[:int(tr_off[cut])], tr_off[:cut], int(tr_base[cut])
        tr_base = tr_base[:cut]
    tr = EncodedBattleFrames(max(tr_rows, 1), device=local, engine=eng)
    tr_out = tr._out()
    d_ts = torch.from_numpy(tr_stream).to(dev)
    d_to = torch.from_numpy(tr_off.view(np.int64)).to(dev)
    d_tb = torch.from_numpy(tr_base.view(np.int32)).to(dev)
    TR_IT = 5
    eng.rows_build_device(d_ts, tr_stream.size, d_to, d_tb, len(tr_off), tr_rows, tr_out)
    torch.cuda.synchronize(dev)
    barrier()
    tr_ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    tr_ev[0].record()
    for _ in range(TR_IT):
        eng.rows_build_device(d_ts, tr_stream.size, d_to, d_tb, len(tr_off), tr_rows, tr_out)
    tr_ev[1].record()
    torch.cuda.synchronize(dev)
    tr_ms = tr_ev[0].elapsed_time(tr_ev[1]) / TR_IT
    assert eng.rows_status() == -1, "a self-play game did not replay to its stored result"
    tr_pin = torch.from_numpy(tr_stream).pin_memory()
    eng.rows_build_host(tr_pin, tr_out, tr_rows)
    t0 = time.perf_counter()
    for _ in range(TR_IT):
        eng.rows_build_host(tr_pin, tr_out, tr_rows)  # index walk + upload + replay + encode + status read-back
    tr_e2e_s = (time.perf_counter() - t0) / TR_IT
    launches += 2 * (2 * TR_IT + 2)
    tr_cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        from oak_ref import Oracle as _O3, oracle_training_rows as _otr
        _cut = int(np.searchsorted(tr_base, 1 << 17, side="right")) - 1  # ~131k rows: about a second of CPU work
        _sub = tr_stream[:int(tr_off[_cut])] if _cut > 0 else tr_stream
        _o3 = _O3()
        t0 = time.perf_counter()
        _r = _otr(_o3, _sub, threads=os.cpu_count() or 1)
        tr_cpu = _r["k"].shape[0] / (time.perf_counter() - t0)
        del _r
    TR_ROW_BYTES = 4 * (2376 + 458 + 12 + 18 + 18 + 3) + 8 * 18 + 2 + 2 + 4  # every output array, per row
    del tr, tr_out

    # ---- secondary: tree search with GPU leaves (BASELINE configs[3]): 4096 root positions sharded over the ranks,
    # host trees + UCB bandits, 64 descents per root in flight, clamped rolls 3/1 as the reference's default_search
    from oak_b200.search import TreeSearch
    SR_ROOTS, SR_P, SR_ITERS = max(1, 4096 // world), 64, 256
    sr_b, sr_d, sr_r = (x[:SR_ROOTS].cpu().numpy() for x in (lock_b, lock_d, lock_r))  # states after 19 playout steps
    sr_live = (sr_r & 15) == 0
    sr_b, sr_d, sr_r = sr_b[sr_live], sr_d[sr_live], sr_r[sr_live]
    sr_seeds = seeds_np[:SR_ROOTS][sr_live]
    ts = TreeSearch(eng, sr_b.shape[0], particles=SR_P)
    ts.set_roots(sr_b, sr_d, sr_r, sr_seeds)
    ts.run(SR_ITERS)  # warm-up: the same search once; the search object keeps its tree arenas, as it does from turn to turn
    ts.set_roots(sr_b, sr_d, sr_r, sr_seeds)
    barrier()
    l0 = eng.launch_count
    t0 = time.perf_counter()
    ts.run(SR_ITERS)
    barrier()
    sr_s = time.perf_counter() - t0
    sr_timing = ts.timing()
    launches += eng.launch_count - l0
    sr_out = ts.output()
    sr_done = float(sr_b.shape[0] * SR_ITERS)
    ts.close()
    sr_cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        import ctypes as _ct
        from concurrent.futures import ThreadPoolExecutor
        from oak_ref import Oracle as _O2
        _f = _O2().lib.oracle_mcts_run
        _f.restype = None
        _f.argtypes = [_ct.c_void_p, _ct.c_void_p, _ct.c_uint8, _ct.c_uint64, _ct.c_uint32, _ct.c_float, _ct.c_uint32, _ct.c_uint32,
                       _ct.c_void_p, _ct.c_void_p, _ct.c_void_p]
        _thr = os.cpu_count() or 1
        _nr, _it = 4 * _thr, 4096

        def _one(i):
            v, w = np.zeros(81, np.uint32), np.zeros(81, np.float64)
            _f(sr_b[i].ctypes.data, sr_d[i].ctypes.data, int(sr_r[i]), int(sr_seeds[i]), _it, 1.0, 3, 1, v.ctypes.data, w.ctypes.data, None)
        t0 = time.perf_counter()
        with ThreadPoolExecutor(_thr) as ex:
            list(ex.map(_one, range(min(_nr, sr_b.shape[0]))))
        sr_cpu = min(_nr, sr_b.shape[0]) * _it / (time.perf_counter() - t0)

    eval_cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        from oak_ref import NetOracle, Oracle as _O
        _orc = _O()
        _net = NetOracle(_orc, blob)
        cb, cd, cr = (x.cpu().numpy() for x in (eb, ed, er))
        _thr = os.cpu_count() or 1
        # positions grouped by the battle they come from (team pair = index & 255): the reference fills its embedding memo
        # once per search root (network.h:48-50) and every evaluation under that root hits it
        _grp = (np.arange(EV) & 255).astype(np.uint32)
        _ord = np.argsort(_grp, kind="stable")
        gb, gd, gr, gg = cb[_ord], cd[_ord], cr[_ord], _grp[_ord]
        _one = slice(0, min(EV, 8192))  # 1-thread sample: 32 groups of 256... (sorted: whole groups)

        def _time(fn):
            t0 = time.perf_counter()
            fn()
            return time.perf_counter() - t0
        eval_cpu = {
            "cache_free_all_cores": EV / _time(lambda: _net.eval_batch(cb, cd, cr, threads=_thr)),
            "cache_free_1_thread": (_one.stop) / _time(lambda: _net.eval_batch(cb[_one], cd[_one], cr[_one], threads=1)),
            "memoised_256_per_root_all_cores": EV / _time(lambda: _net.eval_batch_cached(gb, gd, gr, gg, threads=_thr)),
            "memoised_4096_per_root_all_cores": 16 * EV / _time(lambda: _net.eval_batch_cached(gb, gd, gr, gg, threads=_thr, repeat=16)),
            "memoised_4096_per_root_1_thread": 16 * _one.stop / _time(lambda: _net.eval_batch_cached(gb[_one], gd[_one], gr[_one], gg[_one], threads=1, repeat=16)),
            "cores": _thr,
            "note": "memoised = the reference's algorithm: sparse first layers + PokemonCache (240 keyed embeddings per party slot, "
                    "filled once per root: 2 880 embeddings, the cost of ~290 cache-free evaluations) + ActivePokemonCache map "
                    "(nn/battle/cache.h); N_per_root = evaluations sharing one fill, as a search of N iterations does; cache_free = every embedding recomputed, "
                    "what the GPU kernel does. The reference prints ~85 K evals/s per thread for the memoised path (TUTORIAL.md:418-427)",
        }

    # ---- reduce over ranks: max time, sum of work
    t = torch.tensor([ms_total, e2e_s * 1e3, lock_ms, sr_s * 1e3, tr_ms, tr_e2e_s * 1e3], dtype=torch.float64, device=dev)
    w = torch.tensor([steps_per_pass * args.steps, n * args.steps, e2e_steps * e2e_iters, launches, n, sr_done, tr_rows], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(w, op=dist.ReduceOp.SUM)
    ms_total_max, e2e_ms_max, lock_ms_max, sr_ms_max, tr_ms_max, tr_e2e_ms_max = t.tolist()
    tot_steps, tot_rollouts, tot_e2e_steps, tot_launches, tot_battles, tot_sr, tot_tr_rows = w.tolist()

    if rank == 0:
        pk, pk_src = peaks()
        value = tot_steps / (ms_total_max * 1e-3)
        kern_s = ms_total_max / args.steps * 1e-3  # time per launch over the timed region (launches overlap only at their tails)
        achieved = B_ROLLOUT * n / kern_s / 1e9
        # DRAM traffic of one launch: from the newest `ncu --set full` summary when it is of the bench configuration (2^20
        # playouts per launch), else the round-1 capture of the same configuration; scaled to this n
        iss_ = issue_stats()
        if iss_ and "2^20" in iss_.get("command", "") and iss_.get("dram_bytes"):
            traffic_bytes = int(iss_["dram_bytes"] * n / (1 << 20))
            traffic_src = (f"dram__bytes_read.sum + dram__bytes_write.sum of one launch of 2^20 playouts (ncu --set full, {iss_['file']}, "
                           f"captured {iss_.get('captured')}), scaled to this n")
        else:
            traffic_bytes = int(424_100_608 * n / (1 << 20))
            traffic_src = ("dram__bytes_read.sum + dram__bytes_write.sum = 414.9 + 9.2 MB for one launch of 2^20 playouts (ncu --set full, "
                           "profiles/r1m_rollout_kernel_ncu_summary.txt), scaled to this n")
        cpu = None
        if not args.no_cpu_baseline:
            from oak_ref import Oracle
            threads = os.cpu_count() or 1
            orc_, s, dt = Oracle(), 0, 0.0
            for chunk in range(max(1, args.cpu_sample >> 20)):  # 2^20-playout chunks (412 MB each)
                s_, dt_ = cpu_rollouts(orc_, chunk << 20, min(args.cpu_sample, 1 << 20), threads)
                s, dt = s + s_, dt + dt_
            s1, dt1 = cpu_rollouts(orc_, 0, 1 << 15, 1)
            cpu = {"value": s / dt, "unit": "battle-steps/s", "cores": threads, "kind": "port",
                   "rollouts_per_s": args.cpu_sample / dt,
                   "one_thread": {"battle_steps_per_s": s1 / dt1, "rollouts_per_s": (1 << 15) / dt1, "sample": "32768 playouts"},
                   "sample": f"{args.cpu_sample} playouts ({s} battle-steps, {dt:.1f} s) of the same synthetic workload; "
                             "CPU restatement of libpkmn/Oak (Zig toolchain unavailable)"}
        line = {
            "metric": "battle-steps/sec", "value": value, "unit": "battle-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total_max / args.steps, "higher_is_better": True, "scaling": "weak",
            "legs": {"rollouts_per_s": tot_rollouts / (ms_total_max * 1e-3), "e2e_battle_steps_per_s": tot_e2e_steps / (e2e_ms_max * 1e-3),
                     "evals_per_s": world * EV / (eval_ms * 1e-3), "evals_e2e": world * EV / eval_e2e_s,
                     "search_iters_per_s": tot_sr / (sr_ms_max * 1e-3), "selfplay_updates_per_s": world * SP_G * SP_STEPS / sp_s,
                     "rows_per_s": tot_tr_rows / (tr_ms_max * 1e-3), "lockstep_steps_per_s": tot_battles / (lock_ms_max * 1e-3),
                     "gather_replay_ok": gather_ok},
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "rollouts_per_s": tot_rollouts / (ms_total_max * 1e-3),
            "mean_steps_per_rollout": tot_steps / tot_rollouts,
            "config": {"workload": "configs[1]: 2^20 lockstep random-playout RBY OU sample-team battles per GPU, played to "
                                   "terminal by the fused rollout kernel (choices x2 + pick + update per battle-step)",
                       "battles_per_gpu": n, "parallelism": f"battle-sharded x{world}, no data-path collective",
                       "streams": "steps issued alternately on two CUDA streams: one launch's serial tail overlaps the next launch's body",
                       "l2": f"inputs {n * 393 / 1e6:.0f} MB per pass > 126 MB L2 (no flush needed)"},
            "e2e": {"value": tot_e2e_steps / (e2e_ms_max * 1e-3), "unit": "battle-steps/s",
                    "rollouts_per_s": tot_battles * e2e_iters / (e2e_ms_max * 1e-3),
                    "h2d_bytes_per_step": int(tot_battles * (384 + 8 + 1 + 8)), "d2h_bytes_per_step": int(tot_battles * (1 + 2 + 4)),
                    "iters": e2e_iters, "timing": "wall clock around oakb200_rollout_host calls (pinned host records in, score/turns/steps out; chunked upload pipelined behind the playouts) issued alternately by two host threads on two contexts, so one call's upload and tail overlap the other's playouts"},
            "roofline": {"bound": "hbm", "kernel": "rollout_kernel<128,false,false,SPLIT>", "achieved": achieved, "peak": pk["hbm_gbs"],
                         "unit": "GB/s", "frac": achieved / pk["hbm_gbs"],
                         "traffic": traffic_bytes, "traffic_source": traffic_src,
                         "peak_source": pk_src,
                         "algorithmic_bytes_per_launch": B_ROLLOUT * n, "kernel_ms": ms_total_max / args.steps,
                         "launch_ms_on_own_stream": statistics.mean(kernel_ms), "launch_ms_alone": alone_ms,
                         "note": "fused playouts read each 393-byte record once per ~120 updates: the kernel is bound by SM "
                                 "instruction issue under divergence, not HBM (see profiles/)",
                         "issue": issue_stats()},
            "lockstep_step_kernel": {"ms_per_launch": lock_ms_max, "battles": int(tot_battles),
                                     "battle_steps_per_s": tot_battles / (lock_ms_max * 1e-3),
                                     "achieved_gbs": B_STEP * n / (lock_ms * 1e-3) / 1e9,
                                     "frac_of_hbm_peak": B_STEP * n / (lock_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                                     "live_fraction_at_start": live_before / n},
            "leaf_eval": {"evals_per_s": world * EV / (eval_ms * 1e-3), "batch_per_gpu": EV, "ms_per_batch": eval_ms, "ms_per_batch_one_stream": eval_ms_one,
                          "streams": "batches issued alternately on two CUDA streams (the next launch's CTAs take the SMs that have no fourth tile); one_stream = launches back to back on one stream",
                          "network": "default dims 198/427->128->59/83, 768->64->64, value 32, policy 64->315 (legal rows), relu",
                          "kernel": "eval_tc_kernel (tcgen05.mma kind::f16 on split-bf16 operand pairs, fp32 accumulate in TMEM; "
                                    "one-hot first layers as dense tiles)",
                          "e2e_evals_per_s": world * EV / eval_e2e_s,
                          "e2e_evals_per_s_one_caller": world * EV / eval_e2e_one_s,
                          "e2e_note": "pinned host records in (393 B/eval H2D), value + logits + counts out (78 B/eval D2H), wall clock around oakb200_eval_host (chunked upload / kernel / download on three streams); e2e_evals_per_s = calls issued by two host threads on two contexts (one call's upload under the other's kernels and download), one_caller = calls one after another",
                          "fp32_cuda_core_kernel_evals_per_s": world * EV / (eval_ms_mode[1] * 1e-3),
                          "roofline": {"bound": "tensor", "achieved": F_EVAL_TC * EV / (eval_ms * 1e-3) / 1e12,
                                       "peak": pk["bf16_tflops"], "unit": "TFLOP/s",
                                       "frac": F_EVAL_TC * EV / (eval_ms * 1e-3) / 1e12 / pk["bf16_tflops"], "peak_source": pk_src,
                                       "flops_per_eval_executed": F_EVAL_TC, "flops_per_eval_dense_equivalent": F_EVAL,
                                       "dense_equivalent_tflops": F_EVAL * EV / (eval_ms * 1e-3) / 1e12,
                                       "note": "executed = every tcgen05.mma issued (3 products per split-bf16 GEMM, K and N "
                                               "padding included); dense-equivalent = SURVEY 8d F_eval"},
                          "cpu_evals_per_s": eval_cpu, "cpu_kind": "port (oracle/nn_oracle.hpp)"},
            "tree_search": {"iterations_per_s": tot_sr / (sr_ms_max * 1e-3), "roots_total": int(tot_sr / SR_ITERS),
                            "rank0_ms_per_wave": {"total": round(sr_timing["total_ms"] / max(sr_timing["waves"], 1), 2),
                                                  "host_phases": round(sr_timing["host_ms"] / max(sr_timing["waves"], 1), 2),
                                                  "all_groups_on_device": round(sr_timing["device_wait_ms"] / max(sr_timing["waves"], 1), 2),
                                                  "note": "host_phases = tree work of the thread pool + copies and launches, one group at a "
                                                          "time; the rest of the wave the calling thread polls streams that are still busy"},
                            "descents_in_flight_per_root": SR_P, "iterations_per_root": SR_ITERS,
                            "mean_root_value": float(sr_out["empirical_value"].mean()),
                            "cpu_iterations_per_s": sr_cpu, "cpu_kind": "port (sequential MCTS restatement, all host cores)",
                            "note": "MonteCarlo leaves, joint UCB c=1, clamped rolls 3/1 (mcts.h:133); host trees, one kernel "
                                    "launch per tree level + hidden-variable draw + playouts per wave"},
            "selfplay": {"game_updates_per_s": world * SP_G * SP_STEPS / sp_s, "games_per_gpu": SP_G, "lanes": SP_LANES, "steps": SP_STEPS,
                         "passes_ms": sp_all_s, "rank0_ms": {"generate": sp_phases_timed[0], "flush": sp_phases_timed[1], "barrier": sp_phases_timed[2],
                                      "writer": sp_phases_timed[3]},
                         "stream_bytes_at_writer": int(sp_bytes),
                         "note": "upper bound on recorded updates/s (slots whose game ended wait <= 8 steps for a refill); per "
                                 "step: choices + tensor-core eval + policy sample + frame record + update + finish kernels; "
                                 "drain every 8 steps; every rank DMAs its drains into its own page-locked shared-memory ring, one writer "
                                 "thread on rank 0 reads all rings (no collective, no padding, no rank-0 PCIe funnel)"},
            "training_rows": {"rows_per_s": tot_tr_rows / (tr_ms_max * 1e-3), "rows_per_gpu": int(tr_rows), "games_per_gpu": int(len(tr_off)),
                              "ms_per_build": tr_ms_max, "e2e_rows_per_s": tot_tr_rows / (tr_e2e_ms_max * 1e-3),
                              "e2e_note": "pinned host .battle.data stream in (header walk, H2D), rows left in device memory for the "
                                          "learner, status word back; wall clock around oakb200_rows_build_host",
                              "input": "the games this rank's GPU self-play finished (its own .battle.data stream)",
                              "roofline": {"bound": "hbm", "kernel": "replay_kernel + rows_encode_kernel", "achieved": TR_ROW_BYTES * tr_rows / (tr_ms * 1e-3) / 1e9,
                                           "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": TR_ROW_BYTES * tr_rows / (tr_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                                           "algorithmic_bytes_per_row": TR_ROW_BYTES, "peak_source": pk_src,
                                           "note": "write-only stream: torch's fill of 40 GiB reaches 3.9 TB/s on this pool "
                                                   "(tools/memset_bw.py), the copy peak counts read + write"},
                              "cpu_rows_per_s": tr_cpu, "cpu_kind": "port (oracle_training_rows, all host cores)"},
            "cpu_baseline": cpu, "clocks": clocks, "gpu_launches": int(tot_launches),
        }
        emit(line)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
