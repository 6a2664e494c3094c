"""BASELINE.json configs[0]: the reference's own fixture scene (doc/example_data, real depth-camera geometry: holes,
a workspace mask, 513,688 valid points) through voxel down-sample, collision detection and NMS.  Golden vectors:
tests/golden/scene_example.npz, produced by the reference's utils/data_utils.py and utils/collision_detector.py run
unmodified (tests/golden/make_golden.py scene)."""
import math
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(HERE, "golden", "scene_example.npz")


@pytest.fixture(scope="module")
def scene(b200):
    return b200.synth.fixture_scene(PATH)


def test_fixture_cloud_is_the_references(scene):
    pts, g = scene
    assert pts.shape == (int(g["num_points"]), 3) and pts.dtype == np.float64
    assert (pts[::997] == g["cloud_every_997"]).all()                  # bit-equal to create_point_cloud_from_depth_image
    assert np.allclose(pts.sum(axis=0), g["cloud_sum"], rtol=1e-12)
    assert (pts[:, 2] > 0).all()


def test_oracle_on_the_fixture(scene, oracle_mod, b200):
    """the CPU restatements on real data: voxel count, and detect() == the reference's own detect() (masks, empty
    masks, the five IoU arrays)"""
    pts, g = scene
    vox = oracle_mod.voxel_down_sample(pts, 0.01)
    assert len(vox) == int(g["num_voxels"])
    sample = pts[g["sample_idx"]].astype(np.float32)
    grasps = b200.synth.synthetic_grasps(sample, 1024, int(g["grasp_seed"]))
    assert (oracle_mod.collision_detect(vox, grasps, 0.01, 0.05, 0.01) == g["mask_voxel"]).all()
    assert (oracle_mod.collision_detect(sample.astype(np.float64), grasps, 0.01, 0.05, 0.01) == g["mask_sample"]).all()


@pytest.mark.gpu
def test_detector_on_the_fixture_equals_reference(scene, b200, oracle_mod):
    pts, g = scene
    sample = pts[g["sample_idx"]].astype(np.float32)
    grasps = b200.synth.synthetic_grasps(sample, 1024, int(g["grasp_seed"]))
    gg = b200.GraspGroup(grasps)
    det = b200.ModelFreeCollisionDetector(pts, voxel_size=0.01)
    assert det.scene_points.shape == (int(g["num_voxels"]), 3)
    mask, empty, ious = det.detect(gg, approach_dist=0.05, collision_thresh=0.01, return_empty_grasp=True, return_ious=True)
    assert (mask == g["mask_voxel"]).all() and (empty == g["empty_voxel"]).all()
    assert (np.stack(ious) == g["ious_voxel"]).all()                    # fp64 IoUs from integer counts: bit-equal
    det.scene_points = sample.astype(np.float64)                        # the unvoxelised 20000-point sample
    assert (det.detect(gg, approach_dist=0.05, collision_thresh=0.01) == g["mask_sample"]).all()
    kept = gg[~mask].nms()
    want = oracle_mod.grasp_nms(grasps[~mask])
    assert (kept.grasp_group_array == grasps[~mask][want]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("padded", [False, True])
def test_engine_on_the_fixture_sample(scene, b200, oracle_mod, padded):
    """the whole step on real camera geometry: the 20000-point sample drawn as demo_pointcloud.py:178-186 draws it, and
    a duplicate-padded one (15000 distinct points + 5000 repeats: exact ties for the sampling and the queries)"""
    import torch
    pts, g = scene
    idx = g["sample_idx"]
    if padded:
        rng = np.random.RandomState(1)
        idx = np.concatenate([idx[:15000], rng.choice(idx[:15000], 5000, replace=True)])
    cloud = pts[idx].astype(np.float32)
    net = b200.graspnet.GraspNet(is_training=False)
    b200.synth.fill_state_dict(net, 12)
    net = net.cuda().eval()
    eng = b200.graspnet.GraspNetEngine(net, batch=1, num_point=20000)
    res = eng.infer(cloud[None])[0]
    ep = eng.end_points
    assert (ep["sa1_inds"].cpu().numpy() == oracle_mod.furthest_point_sampling(cloud[None], 2048)).all()
    with torch.no_grad():
        rows = b200.graspnet.pred_decode(net({"point_clouds": torch.from_numpy(cloud[None]).cuda()}))[0].cpu().numpy()
    assert res["nrows"] == len(rows)
    vox = oracle_mod.voxel_down_sample(cloud.astype(np.float64), 0.01)
    mask = oracle_mod.collision_detect(vox, rows, 0.01, 0.05, 0.01)
    assert (res["mask"] == mask).all()
    free = rows[~mask]
    want = free[oracle_mod.grasp_nms(free)]
    assert res["kept"].shape == want.shape and (res["kept"] == want).all()
