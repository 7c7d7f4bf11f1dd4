"""Generates tests/golden/ref_trajectories.npz from oracle/_ref = the reference's OWN node sources
(/root/reference/src/*.cpp) compiled by oracle/build_ref.sh.  Only runnable where /root/reference is
mounted (the build container); the fixture travels.  Per sweep: every pose array of
loam_sweep_result, the feature / map counts, and an order-sensitive checksum of each cloud."""
import importlib
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import checker  # noqa: E402

synth = importlib.import_module("loam_velodyne-1_b200.synth")
# (preset, scene) -> sweeps.  Scene 0 = the BASELINE.json configs at their FULL lengths (200 VLP-16 / 500
# HDL-64 sweeps; 128 rings + IMU: 100); scenes 1..3 = the branch-coverage scenes of synth.cpp (tunnel /
# plain / sparse: isDegenerate, few rows, small corner clouds).
CASES = {(1, 0): 200, (2, 0): 500, (4, 0): 100, (1, 1): 40, (1, 2): 40, (1, 3): 60}
# the same reference sources built against glibc's trig (libloam_ref_libm.so): poses only, for the
# north-star tolerance check (1e-4 rad / 1e-3 m per sweep, end drift <= 0.1 % of the path length)
LIBM_CASES = {(1, 0): 200, (2, 0): 500}
INT_KEYS = ["sr_status", "lo_status", "lm_status", "n_full", "n_sharp", "n_less_sharp", "n_flat", "n_less_flat",
            "n_map_corner", "n_map_surf", "n_stack_corner", "n_stack_surf"]
POSE_KEYS = ["transform", "transform_sum", "transform_aft_mapped", "transform_bef_mapped", "transform_mapped", "imu_trans"]
CLOUDS = [0, 1, 2, 3, 4]          # scanRegistration outputs, every sweep
MAP_CLOUDS = [5, 6, 7, 8, 10, 11, 12, 13]  # on mapping sweeps


def cloud_checksum(c):
    """order-sensitive 64-bit checksum of the float bit patterns"""
    u = np.ascontiguousarray(c, np.float32).view(np.uint32).astype(np.uint64).ravel()
    w = (np.arange(u.size, dtype=np.uint64) * np.uint64(2654435761) + np.uint64(1)) & np.uint64(0xFFFFFFFF)
    return np.uint64((u * w).sum(dtype=np.uint64)) ^ np.uint64(u.size)


def run_case(preset, scene, kind, nsw):
    """one (preset, scene) run of the reference's nodes; returns {name: array}"""
    tag = f"cfg{preset}" + (f"s{scene}" if scene else "")
    S = synth.Synth(preset, scene=scene)
    C = checker.Checker(checker.make_params(preset), kind)
    ints = np.zeros((nsw, len(INT_KEYS)), np.int32)
    poses = np.zeros((nsw, 42), np.float32)
    sums = np.zeros((nsw, 14), np.uint64)
    for k in range(nsw):
        xyz, st = S.sweep(k)
        if S.has_imu:
            for j in range(k * 20, (k + 1) * 20):
                C.imu(*S.imu(j))
        r = C.push(xyz, st)
        d = r.as_dict()
        ints[k] = [d[key] for key in INT_KEYS]
        poses[k] = np.concatenate([d[key] for key in POSE_KEYS])
        if kind != "reference":
            continue
        if r.sr_status == 0:
            for w in CLOUDS:
                sums[k, w] = cloud_checksum(C.cloud(w))
        if r.lm_status == 0:
            for w in MAP_CLOUDS:
                sums[k, w] = cloud_checksum(C.cloud(w))
    C.close()
    if kind == "reference":
        return {f"{tag}_ints": ints, f"{tag}_poses": poses, f"{tag}_sums": sums}
    return {f"{tag}_libm_poses": poses}


def main():
    # one process per case: the reference's nodes keep their state in file-scope globals, and a fresh
    # process is the one way to be sure that no case sees another one's
    import subprocess
    import tempfile
    if len(sys.argv) == 6 and sys.argv[1] == "--case":
        preset, scene, kind, nsw, path = int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], int(sys.argv[5]), None
        np.savez(sys.stdout.buffer, **run_case(preset, scene, kind, nsw))
        return
    checker.build()
    jobs = [(p, s, "reference", n) for (p, s), n in CASES.items()] + [(p, s, "reference-libm", n) for (p, s), n in LIBM_CASES.items()]
    procs = []
    for p, s, kind, n in jobs:
        f = tempfile.TemporaryFile()
        procs.append((subprocess.Popen([sys.executable, os.path.abspath(__file__), "--case", str(p), str(s), kind, str(n)], stdout=f), f, (p, s, kind, n)))
    out = {}
    for pr, f, job in procs:
        assert pr.wait() == 0, job
        f.seek(0)
        with np.load(f) as z:
            out.update({k: z[k] for k in z.files})
        print(job, "done", flush=True)
    out["int_keys"] = np.array(INT_KEYS)
    out["pose_keys"] = np.array(POSE_KEYS)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_trajectories.npz"), **out)
    print("written", {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    main()
