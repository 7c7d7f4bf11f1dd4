"""GPU (-m gpu): the 64-sequence configuration (BASELINE.json configs[4]) -- several pipelines on one
GPU, fed from host threads, with the cooperative LM kernels on a sub-grid so that they co-reside."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

KEYS = ("transform", "transform_sum", "transform_aft_mapped", "transform_mapped")


def _bits(r):
    return b"".join(bytes(getattr(r, k)) for k in KEYS) + bytes([r.lo_iters, r.lm_iters])


@pytest.mark.parametrize("ctas", [37, 16])
def test_sub_grid_cooperative_kernels_keep_parity(pkg, checker, synth, ctas):
    S = synth.Synth(1)
    P = pkg.make_params(1, system_delay=0)
    P.coop_ctas = ctas
    G = pkg.Pipeline(P)
    C = checker.Checker(checker.make_params(1), "port")
    for k in range(30):
        xyz, st = S.sweep(k)
        g, r = G.push(xyz, st), C.push(xyz, st)
        assert (g.lo_iters, g.lm_iters, g.lo_status, g.lm_status) == (r.lo_iters, r.lm_iters, r.lo_status, r.lm_status), k
        for key in KEYS:
            assert bytes(getattr(g, key)) == bytes(getattr(r, key)), (k, key)


def test_concurrent_pipelines_from_host_threads_equal_sequential_runs(pkg, synth):
    n_seq, nsw = 6, 24
    P = pkg.make_params(1, system_delay=0)
    P.map_arena_points = 1 << 20
    P.coop_ctas = 37
    data = {}
    for s in range(n_seq):
        S = synth.Synth(1, seed=1000 + s)
        data[s] = [S.sweep(k) for k in range(nsw)]
    ref = {}
    for s in range(n_seq):
        G = pkg.Pipeline(P)
        ref[s] = [_bits(G.push(x, st)) for x, st in data[s]]
        del G
    pipes = {s: pkg.Pipeline(P) for s in range(n_seq)}
    out = {s: [] for s in range(n_seq)}
    errs = []

    def worker(group):
        try:
            for k in range(nsw):
                for s in group:
                    pipes[s].push_async(data[s][k][0], data[s][k][1])
            for s in group:
                out[s] = [_bits(r) for r in pipes[s].collect()]
        except Exception as e:  # surfaced below
            errs.append(e)

    ths = [threading.Thread(target=worker, args=([s for s in range(n_seq) if s % 3 == t],)) for t in range(3)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errs, errs
    for s in range(n_seq):
        assert out[s] == ref[s], s
    assert len({tuple(v) for v in ref.values()}) == n_seq   # the sequences really differ


def test_all_64_sequences_on_sub_grids_equal_the_oracle(pkg, checker, synth):
    """BASELINE.json configs[4] as bench.py --workload seq64 runs it: the 64 seeds (1000..1063), 20 sweeps
    each, one pipeline per sequence on ONE GPU with the cooperative LM kernels on 37-CTA sub-grids, fed by 8
    host threads -- every pose and iteration count of every sequence against the oracle (run on host threads
    beside it; ctypes releases the GIL)"""
    n_seq, nsw, n_threads = 64, 20, 8
    P = pkg.make_params(1, system_delay=0)
    P.map_arena_points = 1 << 20
    P.coop_ctas = 37
    data = {}
    for s in range(n_seq):
        S = synth.Synth(1, seed=1000 + s, threads=2)
        data[s] = [S.sweep(k) for k in range(nsw)]
    pipes = {s: pkg.Pipeline(P) for s in range(n_seq)}
    got = {s: [] for s in range(n_seq)}
    want = {s: [] for s in range(n_seq)}
    errs = []

    def gpu_worker(group):
        try:
            for k in range(nsw):
                for s in group:
                    pipes[s].push_async(data[s][k][0], data[s][k][1])
            for s in group:
                got[s] = [_bits(r) for r in pipes[s].collect()]
        except Exception as e:
            errs.append(e)

    def cpu_worker(group):
        try:
            for s in group:
                C = checker.Checker(checker.make_params(1), "port")
                want[s] = [_bits(C.push(x, st)) for x, st in data[s]]
                C.close()
        except Exception as e:
            errs.append(e)

    ths = [threading.Thread(target=gpu_worker, args=([s for s in range(n_seq) if s % n_threads == t],)) for t in range(n_threads)]
    ths += [threading.Thread(target=cpu_worker, args=([s for s in range(n_seq) if s % n_threads == t],)) for t in range(n_threads)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errs, errs
    for s in range(n_seq):
        assert len(got[s]) == nsw and got[s] == want[s], s


@pytest.mark.parametrize("system_delay", [0, 2])
def test_batched_scan_registration_equals_per_sequence_pipelines(pkg, synth, system_delay):
    """loam_pipeline_batch_*: the scanRegistration passes of many sequences as one launch chain (blockIdx.z =
    sequence).  Ragged point counts (dropped returns), a sequence with NaNs, systemDelay swallowing the first
    sweeps: every pose, count and feature cloud of every sequence equals what its own pipeline produces"""
    n_seq, nsw = 7, 18
    P = pkg.make_params(1, system_delay=system_delay)
    P.map_arena_points = 1 << 20
    P.coop_ctas = 37
    data = {}
    rng = np.random.default_rng(3)
    for s in range(n_seq):
        S = synth.Synth(1, seed=2000 + s)
        sw = []
        for k in range(nsw):
            x, st = S.sweep(k)
            if s % 3 == 1:
                x = x[: len(x) - 800 * (k % 4)]          # ragged: fewer points on some sweeps
            if s == 2:
                x = x.copy()
                x[rng.choice(len(x), 200, replace=False), 0] = np.nan
            sw.append((x, st))
        data[s] = sw
    ref = {}
    clouds = {}
    for s in range(n_seq):
        G = pkg.Pipeline(P)
        ref[s] = [_bits(G.push(x, st)) for x, st in data[s]]
        clouds[s] = [G.sr.cloud(w).tobytes() for w in (0, 1, 2, 3, 4)]
        del G
    l0 = pkg.kernel_launch_count()
    B = pkg.PipelineBatch(P, n_seq)
    for k in range(nsw):
        B.push_async([data[s][k][0] for s in range(n_seq)], [data[s][k][1] for s in range(n_seq)])
    got = {s: [_bits(r) for r in B.pipes[s].collect()] for s in range(n_seq)}
    launches = pkg.kernel_launch_count() - l0
    for s in range(n_seq):
        assert got[s] == ref[s], s
        assert [B.pipes[s].sr.cloud(w).tobytes() for w in (0, 1, 2, 3, 4)] == clouds[s], s
    # the registration chain is nine launches per step whatever the number of sequences
    assert launches < nsw * (9 + n_seq * 24)
