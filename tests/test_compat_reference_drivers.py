"""CPU, build container only (skipped where /root/reference is absent): the reference's OWN
driver modules import and construct on top of the compat modules (seam 2) —
visual_slam.py / map.py / tracked_*.py / main*.py are used unmodified from /root/reference/src."""
import importlib
import os
import sys

import numpy as np
import pytest

REF = "/root/reference/src"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference sources not on this box")

_NAMES = ["initials", "image_pair", "frame", "frame_generator", "image_and_keypoints",
          "triangulate_points_from_images", "visual_slam", "map", "tracked_camera", "tracked_point",
          "observation", "main", "main_triangulate"]


@pytest.fixture()
def compat_path():
    from visual_odometry_b200 import compat
    saved_path, saved_mods = list(sys.path), {k: sys.modules.get(k) for k in _NAMES}
    for k in _NAMES:
        sys.modules.pop(k, None)
    sys.dont_write_bytecode = True
    compat.install(reference_src=REF)
    yield
    sys.path[:] = saved_path
    for k in _NAMES:
        sys.modules.pop(k, None)
        if saved_mods[k] is not None:
            sys.modules[k] = saved_mods[k]


def test_reference_visual_slam_runs_on_compat_modules(compat_path):
    vs = importlib.import_module("visual_slam")            # the reference's file, unmodified
    assert vs.__file__.startswith(REF)
    import image_pair, frame_generator, initials            # ours, shadowing the reference's
    assert "visual_odometry_b200" in image_pair.ImagePair.__module__
    assert vs.ImagePair is image_pair.ImagePair and vs.FrameGenerator is frame_generator.FrameGenerator
    slam = vs.VisualSlam("/nonexistent")
    slam.set_camera_matrix()
    # the committed test.g2o pins f, cx, cy of the live intrinsics
    assert np.allclose([slam.camera_matrix[0, 0], slam.camera_matrix[0, 2], slam.camera_matrix[1, 2]],
                       [802.832, 565.427, 240.124], atol=1e-3)
    assert vs.cv2 is initials.cv2 and hasattr(vs.cv2, "solvePnPRansac")
    # the matcher VisualSlam.__init__ builds (src/visual_slam.py:19) is the B200 one, and so is the
    # ORB detector of the commented-out line 16 / of ImageAndKeypoints and the live SIFT detector (:17)
    from visual_odometry_b200.matcher import B200Matcher, NORM_L2
    from visual_odometry_b200.orb import OrbDetector
    assert isinstance(slam.bf, B200Matcher) and slam.bf.normType == NORM_L2 and slam.bf.crossCheck
    det = vs.cv2.ORB_create(nfeatures=700)
    assert isinstance(det, OrbDetector) and det.nfeatures == 700 and det.nlevels == 8
    with pytest.raises(NotImplementedError):
        vs.cv2.ORB_create(WTA_K=3)
    from visual_odometry_b200.sift import SiftDetector
    assert isinstance(slam.detector, SiftDetector)
    assert vs.cv2.SIFT_create(nfeatures=500).nfeatures == 500
    with pytest.raises(NotImplementedError):
        vs.cv2.SIFT_create(contrastThreshold=0.02)
    # the kNN matcher of src/feature_detection.py:20 (FLANN there) is the exact brute-force one here
    knn = vs.cv2.DescriptorMatcher_create(vs.cv2.DESCRIPTOR_MATCHER_FLANNBASED)
    assert isinstance(knn, B200Matcher) and knn.normType == NORM_L2 and not knn.crossCheck
    assert vs.cv2.DescriptorMatcher_create("BruteForce-Hamming").normType != NORM_L2
    # namedtuples keep the reference's field order
    assert initials.Match._fields == ('featureid1', 'featureid2', 'keypoint1', 'keypoint2',
                                      'descriptor1', 'descriptor2', 'distance', 'color')
    assert initials.Match3D._fields[-1] == 'point'
    assert initials.MatchWithMap._fields[2:4] == ('imagecoord', 'mapcoord')


def test_reference_entry_points_import(compat_path):
    mt = importlib.import_module("main_triangulate")        # reference file, unmodified
    import triangulate_points_from_images as tp
    assert mt.TriangulatePointsFromTwoImages is tp.TriangulatePointsFromTwoImages
    assert "visual_odometry_b200" in tp.TriangulatePointsFromTwoImages.__module__
    from image_pair import ImagePair
    legacy = ImagePair("ORB")                                # the shape main_triangulate.py needs
    assert hasattr(legacy, "standard_pipeline") and hasattr(legacy, "set_images")


def test_install_can_patch_the_reference_map(compat_path):
    """compat.install(patch_map=True): the reference's own Map class gets the GPU optimize_map and
    reprojection methods (binding only: no GPU call here)."""
    from visual_odometry_b200 import compat, ba, map_ops
    sys.modules.pop("map", None)
    compat.install(reference_src=REF, patch_map=True)
    ref_map = importlib.import_module("map")
    assert ref_map.__file__.startswith(REF)
    assert ref_map.Map.optimize_map is ba.optimize_map
    assert ref_map.Map.calculate_reprojection_error is map_ops.calculate_reprojection_error
    m = ref_map.Map()                         # the rest of the class is the reference's own
    assert m.increment_id() == 0 and m.increment_id() == 1


def test_reference_run_loop_returns_when_headless(compat_path, tmp_path, monkeypatch):
    """The reference's UNMODIFIED VisualSlam.run() (src/visual_slam.py:338-374) ends in
    `while True: k = cv2.waitKey(100); if k == ord('q'): break`.  With the headless cv2 namespace
    it must come back: every frame is read and processed, then the idle loop is released."""
    import time
    import cv2 as real_cv2
    vs = importlib.import_module("visual_slam")
    rng = np.random.default_rng(0)
    for k in range(3):
        real_cv2.imwrite(str(tmp_path / f"f{k:02d}.jpg"), rng.integers(0, 255, (48, 64, 3), dtype=np.uint8))
    slam = vs.VisualSlam(str(tmp_path))
    slam.set_camera_matrix()
    seen = []

    def fake_process_frame(img):            # stands for make_frame + match_current_and_previous_frame
        seen.append(img.shape)
        if len(seen) > 1:
            vs.cv2.imshow("matches", img)   # visual_slam.py:304
            vs.cv2.waitKey(100)             # visual_slam.py:312
        return img

    monkeypatch.setattr(slam, "process_frame", fake_process_frame)
    monkeypatch.setattr(slam.map, "show_map_statistics", lambda: None)
    monkeypatch.setenv("VO_B200_HEADLESS_QUIT", "5")
    t0 = time.perf_counter()
    slam.run()
    assert len(seen) == 3 and time.perf_counter() - t0 < 5.0
    # a frame loop with more frames than the idle limit is not cut short: imread / imshow reset it
    for _ in range(12):
        vs.cv2.imread(str(tmp_path / "f00.jpg"))
        assert vs.cv2.waitKey(400000) == -1


def test_patched_map_statistics_on_an_empty_map_like_the_reference(compat_path):
    """VisualSlam.run prints map statistics after frame 0, when the map is empty and its
    camera_matrix is still None (src/visual_slam.py:354, src/map.py:236-239): the reference's loops
    over no observations return 0 and touch nothing; the GPU drop-ins must do the same."""
    from visual_odometry_b200 import compat
    compat.install(reference_src=REF, patch_map=True)
    ref_map = importlib.import_module("map")
    m = ref_map.Map()
    assert m.camera_matrix is None
    assert m.calculate_reprojection_error() == 0
    m.remove_observations_with_reprojection_errors_above_threshold()
    m.show_map_statistics()
