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
CASES = {1: 100, 2: 30, 4: 16}
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


def main():
    checker.build()
    out = {}
    for preset, nsw in CASES.items():
        S = synth.Synth(preset)
        C = checker.Checker(checker.make_params(preset), "reference")
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
            if r.sr_status == 0:
                for w in CLOUDS:
                    sums[k, w] = cloud_checksum(C.cloud(w))
            if r.lm_status == 0:
                for w in MAP_CLOUDS:
                    sums[k, w] = cloud_checksum(C.cloud(w))
        out[f"cfg{preset}_ints"], out[f"cfg{preset}_poses"], out[f"cfg{preset}_sums"] = ints, poses, sums
        C.close()
    out["int_keys"] = np.array(INT_KEYS)
    out["pose_keys"] = np.array(POSE_KEYS)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_trajectories.npz"), **out)
    print("written", {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    main()
