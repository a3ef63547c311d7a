"""Golden vectors from the REFERENCE'S OWN bundle-adjustment code: oracle/_ref/libref_ba.so is
/root/reference/src/bundle_adjuster.cpp + reprojection_factor.cpp compiled where they lie
(oracle/build_ref.py: stand-in Eigen / OpenCV headers and a recording, non-solving Ceres - read
its header for what that does and does not pin). This script drives it and stores inputs and
outputs in tests/golden/ba_reference.npz so that the tests can replay them on machines without
the reference tree (the GPU box). Run in the build container:

    python tests/golden/make_golden_ba.py

Stored:
  * `slide_*`  a 12-keyframe run through a window of 5 (src/vo_node.cpp:36): tracked ids carried
               over, new features added, the window sliding from keyframe 6 on - new_ids per
               keyframe, the problem's block counts, the world points of every id at the end;
  * `b6_*`     the truncation branch (tracked + new > 400, src/bundle_adjuster.cpp:87-92): what
               the reference leaves in Keyframe::new_ids there;
  * `win_*`    every residual block of the final window evaluated by the reference's
               ReprojectionFactor::Evaluate (parameters, observation, residual, both Jacobians);
  * `fac_*`    256 single ReprojectionFactor evaluations on random poses (non-unit quaternions
               too), points and observations.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))

from oracle import ref_ba  # noqa: E402

FOCAL, CX, CY, BASELINE = 718.856, 607.1928, 185.2157, 0.537     # config/kitti00.yaml
WINDOW = 5


def _pose(rng, k):
    q = np.array([1.0, 0.0, 0.0, 0.0]) + rng.normal(scale=0.02, size=4)
    q /= np.linalg.norm(q)
    t = np.array([0.05 * k, -0.01 * k, -0.6 * k]) + rng.normal(scale=0.01, size=3)
    return np.concatenate([q, t]).astype(np.float32)


def _scene(rng, n):
    """n world points in front of the camera and where a camera at `pose` sees them"""
    pts = np.stack([rng.uniform(-12, 12, n), rng.uniform(-3, 3, n), rng.uniform(8, 40, n)], 1)
    return pts.astype(np.float32)


def _project(pose, pts):
    q, t = pose[:4].astype(np.float64), pose[4:].astype(np.float64)
    w, x, y, z = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    g = pts.astype(np.float64) @ R.T + t
    return np.stack([FOCAL * g[:, 0] / g[:, 2] + CX, FOCAL * g[:, 1] / g[:, 2] + CY], 1).astype(np.float32)


def sliding_run(out):
    rng = np.random.default_rng(20261105)
    ba = ref_ba.RefBundleAdjuster(WINDOW, FOCAL, CX, CY, BASELINE)
    live_ids = np.zeros(0, np.uint64)
    world = {}
    n_kf = 12
    tracked_all, new3d_all, newids_all, stats_all, counts = [], [], [], [], []
    for k in range(n_kf):
        pose = _pose(rng, k)
        keep = live_ids[rng.random(live_ids.shape[0]) < 0.72]
        n_new = int(rng.integers(25, 90))
        n_new = min(n_new, ba.max_features - keep.shape[0])            # stay out of the B6 branch
        new3d = _scene(rng, n_new)
        kept_pts = np.array([world[int(i)] for i in keep], np.float32).reshape(-1, 3)
        t2d = _project(pose, kept_pts) + rng.normal(scale=0.3, size=(keep.shape[0], 2)).astype(np.float32)
        n2d = _project(pose, new3d)
        new_ids, after = ba.add_keyframe(pose, keep, t2d, n2d, new3d)
        assert after == n_new and new_ids.shape[0] == n_new
        ba.bundle_adjust()
        assert np.array_equal(ba.last_pose(), pose)                    # no solver: pose untouched
        for i, p in zip(new_ids, new3d):
            world[int(i)] = p
        st = ba.stats()
        assert st["api_errors"] == 0
        tracked_all.append(keep)
        new3d_all.append(new3d)
        newids_all.append(new_ids)
        stats_all.append([st["residual_blocks"], st["parameter_blocks"], st["poses"], st["constant_poses"],
                          st["dropped_with_parameter"]])
        counts.append([keep.shape[0], n_new])
        live_ids = np.concatenate([keep, new_ids])
    all_ids = np.array(sorted(world), np.uint64)
    out["slide_window"] = np.int64(WINDOW)
    out["slide_counts"] = np.array(counts, np.int64)
    out["slide_tracked_ids"] = np.concatenate(tracked_all)
    out["slide_new_3d"] = np.concatenate(new3d_all)
    out["slide_new_ids"] = np.concatenate(newids_all)
    out["slide_stats"] = np.array(stats_all, np.int64)
    out["slide_all_ids"] = all_ids
    out["slide_world_points"] = ba.world_points(all_ids)
    win = ba.evaluate_window()
    for key, val in win.items():
        out["win_" + key] = val
    ba.close()


def truncation_run(out):
    rng = np.random.default_rng(7)
    ba = ref_ba.RefBundleAdjuster(WINDOW, FOCAL, CX, CY, BASELINE)
    pose = _pose(rng, 0)
    first3d = _scene(rng, 380)
    ids, _ = ba.add_keyframe(pose, [], np.zeros((0, 2)), _project(pose, first3d), first3d)
    keep = ids[:350]
    new3d = _scene(rng, 100)
    new_ids, after = ba.add_keyframe(pose, keep, _project(pose, first3d[:350]), _project(pose, new3d), new3d)
    out["b6_first_ids"] = ids
    out["b6_tracked_ids"] = keep
    out["b6_new_3d"] = new3d
    out["b6_new_ids"] = new_ids                      # as the reference leaves them
    out["b6_new_2d_after"] = np.int64(after)
    out["b6_world_points_of_kept"] = ba.world_points(new_ids[new_ids.shape[0] - after:])
    ba.close()


def factor_run(out):
    rng = np.random.default_rng(11)
    n = 256
    poses, points, obs = np.zeros((n, 7)), np.zeros((n, 3)), np.zeros((n, 2))
    resid, jp, jx = np.zeros((n, 2)), np.zeros((n, 2, 7)), np.zeros((n, 2, 3))
    for i in range(n):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        if i % 3:
            q *= rng.uniform(0.6, 1.6)               # Ceres hands over non-unit quaternions mid-solve
        poses[i] = np.concatenate([q, rng.normal(scale=2.0, size=3)])
        points[i] = rng.normal(scale=4.0, size=3) + np.array([0, 0, 15.0])
        obs[i] = rng.uniform(0, 1200, 2)
        resid[i], jp[i], jx[i] = ref_ba.factor_evaluate(obs[i], FOCAL, CX, CY, poses[i], points[i])
    out.update(fac_pose=poses, fac_point=points, fac_obs=obs, fac_resid=resid, fac_jac_pose=jp,
               fac_jac_point=jx, fac_camera=np.array([FOCAL, CX, CY]))


def main():
    out = {}
    sliding_run(out)
    truncation_run(out)
    factor_run(out)
    path = os.path.join(HERE, "ba_reference.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
