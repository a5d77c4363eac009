"""Seam 2 of the drop-in boundary: `ImagePair` with the reference's surface, on libvo_b200.

Mirrors reference src/image_pair.py:222-354 (live class) and the legacy zero-argument API kept
there as a commented block (:46-220) that src/triangulate_points_from_images.py:29-37 still
calls.  Same method names, arguments, attributes, return values and error behaviour; every cv2
*computation* (BFMatcher.match, findEssentialMat, recoverPose, triangulatePoints) is replaced by
a CUDA kernel call.  cv2 is used only for its exception type and for drawing.
"""
import math

import numpy as np
import torch

from . import _lib
from . import geometry as G
from .initials_types import Match, Match3D
from .matcher import B200Matcher, NORM_HAMMING, NORM_L2

try:  # exception type compatible with the caller's `except` clauses
    import cv2 as _cv2
    CvError = _cv2.error
except Exception:  # pragma: no cover
    _cv2 = None

    class CvError(Exception):
        pass


def isRotationMatrix(R):
    """reference src/image_pair.py:12-18"""
    Rt = np.transpose(R)
    shouldBeIdentity = np.dot(Rt, R)
    I = np.identity(3, dtype=R.dtype)
    n = np.linalg.norm(I - shouldBeIdentity)
    return n < 1e-6


def rotationMatrixToEulerAngles(R):
    """reference src/image_pair.py:21-43 (x and z swapped w.r.t. MATLAB)."""
    assert (isRotationMatrix(R))
    sy = math.sqrt(R[0, 0] * R[0, 0] + R[1, 0] * R[1, 0])
    singular = sy < 1e-6
    if not singular:
        x = math.atan2(R[2, 1], R[2, 2])
        y = math.atan2(-R[2, 0], sy)
        z = math.atan2(R[1, 0], R[0, 0])
    else:
        x = math.atan2(-R[1, 2], R[1, 1])
        y = math.atan2(-R[2, 0], sy)
        z = 0
    return np.array([x, y, z])


def _cuda(a, dtype=torch.float64):
    return torch.from_numpy(np.ascontiguousarray(a)).to(device="cuda", dtype=dtype)


# ---------------------------------------------------------------------- single-pair kernels
def find_essential_mat(points1, points2, camera_matrix, prob=0.99, threshold=1.0, max_iters=1000):
    """cv2.findEssentialMat(p1, p2, K, FM_RANSAC, prob, threshold) -> (E [3,3] f64, mask [M,1] u8)."""
    p1 = np.asarray(points1, np.float64).reshape(-1, 2)
    p2 = np.asarray(points2, np.float64).reshape(-1, 2)
    if len(p1) < 5 or len(p1) != len(p2):
        raise CvError("findEssentialMat: needs >= 5 point pairs of equal count (npoints >= 0 && "
                      "points2.checkVector(2) == npoints)")
    pts = _cuda(np.concatenate([p1, p2], 1)[None])
    m = torch.tensor([len(p1)], dtype=torch.int32, device="cuda")
    out = G.find_essential_batch(pts, m, _cuda(camera_matrix), prob, threshold, max_iters)
    if int(out["status"][0]) != 0:
        return None, np.zeros((len(p1), 1), np.uint8)
    return out["E"][0].cpu().numpy(), out["mask"][0, :len(p1)].cpu().numpy().reshape(-1, 1)


def recover_pose(E, points1, points2, camera_matrix, distance_thresh=50.0):
    """cv2.recoverPose(E, p1, p2, K) -> (retval, R [3,3], t [3,1], mask [M,1] u8 {0,255})."""
    p1 = np.asarray(points1, np.float64).reshape(-1, 2)
    p2 = np.asarray(points2, np.float64).reshape(-1, 2)
    E = np.asarray(E, np.float64)
    if E.shape != (3, 3):
        raise CvError("recoverPose: E must be 3x3")
    n = len(p1)
    pts = _cuda(np.concatenate([p1, p2], 1).reshape(1, max(n, 1), 4) if n else np.zeros((1, 1, 4)))
    m = torch.tensor([n], dtype=torch.int32, device="cuda")
    out = G.recover_pose_batch(_cuda(E[None]), pts, m, None, _cuda(camera_matrix),
                               distance_thresh, want_mask=True)
    return (int(out["n_good"][0]), out["R"][0].cpu().numpy(),
            out["t"][0].cpu().numpy().reshape(3, 1),
            out["pose_mask"][0, :n].cpu().numpy().reshape(-1, 1))


def triangulate_points(P1, P2, points1_2xN, points2_2xN):
    """cv2.triangulatePoints(P1, P2, x1 [2,N], x2 [2,N]) -> X [4,N] (dtype follows the points)."""
    x1 = np.asarray(points1_2xN)
    x2 = np.asarray(points2_2xN)
    out_dtype = np.float32 if x1.dtype == np.float32 else np.float64
    X = G.triangulate(_cuda(np.asarray(P1, np.float64)), _cuda(np.asarray(P2, np.float64)),
                      _cuda(x1.T.astype(np.float64)), _cuda(x2.T.astype(np.float64)))
    return X.cpu().numpy().T.astype(out_dtype)


# The reference prints every relative pose and every reconstructed point cloud
# (src/image_pair.py:313-314, 352-354); kept by default (drop-in), switched off by
# VO_B200_QUIET=1 or `image_pair.VERBOSE = False` (the array formatting alone costs ~1 ms per pair)
import os as _os
VERBOSE = _os.environ.get("VO_B200_QUIET", "") in ("", "0")


def _say(*a):
    if VERBOSE:
        print(*a)


class ImagePair():
    """
    Class for working with image pairs (reference src/image_pair.py:222).
    Accepts both constructor shapes: (frame1, frame2, matcher, camera_matrix) — live — and
    (detector_name) — legacy, src/image_pair.py:47-53.
    """

    def __init__(self, frame1, frame2=None, matcher=None, camera_matrix=None):
        if isinstance(frame1, str) and frame2 is None:
            # legacy: ImagePair("ORB") builds its own cross-check matcher
            self.bf = B200Matcher(NORM_HAMMING if frame1 == "ORB" else NORM_L2, crossCheck=True)
            self._legacy = True
            return
        self._legacy = False
        self.frame1 = frame1
        self.frame2 = frame2
        self.matcher = matcher
        self.camera_matrix = camera_matrix

    # ------------------------------------------------------------------ live API (:233-354)
    def match_features(self):
        dmatches = self.matcher.match(self.frame1.descriptors, self.frame2.descriptors)
        feats1, feats2 = self.frame1.features, self.frame2.features
        # colour is unseeded random in the reference too (:250) — excluded from parity; drawn in
        # one call instead of once per match
        colors = np.random.random((len(dmatches), 3))
        raw = []
        for k, dm in enumerate(dmatches):
            a, b = feats1[dm.queryIdx], feats2[dm.trainIdx]
            raw.append(Match(a.feature_id, b.feature_id, a.keypoint.pt, b.keypoint.pt,
                             a.descriptor, b.descriptor, dm.distance, colors[k]))
        self.raw_matches = raw
        # the reference's crude distance gate (src/image_pair.py:254-258; a no-op in practice)
        self.filtered_matches = [mt for mt in self.raw_matches if mt.distance < 1130]
        self._points_cache = []
        p1, p2 = getattr(self.frame1, "points", None), getattr(self.frame2, "points", None)
        if p1 is not None and p2 is not None and len(self.filtered_matches) == len(raw):
            qi = np.fromiter((dm.queryIdx for dm in dmatches), np.int64, len(dmatches))
            ti = np.fromiter((dm.trainIdx for dm in dmatches), np.int64, len(dmatches))
            self._remember_points(self.filtered_matches, p1[qi].astype(np.float64),
                                  p2[ti].astype(np.float64))

    # pixel coordinates of match lists this object produced itself (identity-keyed, so a list
    # built or edited by the caller always takes the generic path below)
    def _remember_points(self, matches, pa, pb):
        cache = getattr(self, "_points_cache", None)
        if cache is None:
            cache = self._points_cache = []
        cache.append((matches, len(matches), np.ascontiguousarray(pa), np.ascontiguousarray(pb)))
        del cache[:-4]

    def visualize_matches(self, matches):
        """side-by-side canvas with one line per match (reference :260-273)"""
        width = self.frame1.image.shape[1]
        canvas = np.concatenate((self.frame1.image, self.frame2.image), axis=1)
        for mt in matches:
            (u1, v1), (u2, v2) = mt.keypoint1, mt.keypoint2
            canvas = _cv2.line(canvas, (int(u1), int(v1)), (int(u2 + width), int(v2)),
                               list(mt.color * 256), 1)
        return canvas

    def get_image_points(self, matches):
        """two float64 [M,2] arrays of pixel coordinates (reference :294-299)"""
        for lst, n, pa, pb in getattr(self, "_points_cache", ()):
            if lst is matches and n == len(matches):
                return pa, pb
        both = np.array([tuple(mt.keypoint1) + tuple(mt.keypoint2) for mt in matches],
                        dtype=np.float64).reshape(-1, 4)
        return np.ascontiguousarray(both[:, :2]), np.ascontiguousarray(both[:, 2:])

    def determine_essential_matrix(self, matches=None):
        if self._legacy:
            return self._legacy_determine_essential_matrix()
        pa, pb = self.get_image_points(matches)
        # confidence 0.99, threshold 1 px: the reference's constants (:278-279)
        self.essential_matrix, mask = find_essential_mat(pa, pb, self.camera_matrix, 0.99, 1)
        keep = mask.ravel() == 1
        inliers = [mt for mt, inlier in zip(matches, keep) if inlier]
        self._remember_points(inliers, pa[keep], pb[keep])
        return inliers

    def estimate_camera_movement(self, matches=None):
        if self._legacy:
            return self._legacy_estimate_camera_movement()
        pa, pb = self.get_image_points(matches)
        _, self.R, self.t, _ = recover_pose(self.essential_matrix, pa, pb, self.camera_matrix)
        pose = np.eye(4)
        pose[:3, :3], pose[:3, 3] = self.R, self.t.ravel()
        self.relative_pose = pose
        _say("relative movement in image pair")
        _say(self.relative_pose)

    def reconstruct_3d_points(self, matches=None,
                              first_projection_matrix=None,
                              second_projection_matrix=None):
        if self._legacy:
            return self._legacy_reconstruct_3d_points()
        K = self.camera_matrix
        # defaults: K[I|0] and K[R^T | -R^T t]; overrides are 3x4 poses multiplied by K (:319-328).
        # NB the reference pairs `projection_matrix` with the frame-1 points — kept as is.
        first = np.eye(3, 4) if first_projection_matrix is None else first_projection_matrix
        second = (np.hstack((self.R.T, -self.R.T @ self.t))
                  if second_projection_matrix is None else second_projection_matrix)
        self.null_projection_matrix = K @ first
        self.projection_matrix = K @ second
        pa, pb = self.get_image_points(matches)
        X = triangulate_points(self.projection_matrix, self.null_projection_matrix, pa.T, pb.T)
        X /= X[3, :]  # unit homogeneous part (:339)
        self.points3d_reconstr = X
        self.matches_with_3d_information = [
            Match3D(*mt, (X[0, k], X[1, k], X[2, k])) for k, mt in enumerate(matches)]
        _say("Reconstructed points")
        if VERBOSE:
            _say(self.points3d_reconstr.transpose().shape)
            _say(self.points3d_reconstr.transpose())

    # ------------------------------------------------------------------ legacy API (:46-220)
    def set_images(self, img1, img2):
        self.image1 = img1
        self.image2 = img2

    def match_detected_keypoints(self):
        """legacy :62-80 — float32 point arrays, match indices, per-keypoint colours"""
        self.matches = self.bf.match(self.image1.descriptors, self.image2.descriptors)
        kp1, kp2 = self.image1.keypoints, self.image2.keypoints
        self.points1_1to2 = np.float32([kp1[m.queryIdx].pt for m in self.matches]).reshape(-1, 2)
        self.points2_1to2 = np.float32([kp2[m.trainIdx].pt for m in self.matches]).reshape(-1, 2)
        self.kp_colors = np.uint8([self.image1.kp_colors[m.queryIdx] for m in self.matches])
        self.match_indices_1to2 = np.arange(len(self.matches), dtype=np.int32)

    def _legacy_determine_essential_matrix(self):
        self.essential_matrix_1to2, mask = find_essential_mat(
            self.points1_1to2, self.points2_1to2, self.image1.cameraMatrix, 0.99, 1)
        keep = mask.ravel() == 1
        self.points1_1to2 = self.points1_1to2[keep]
        self.points2_1to2 = self.points2_1to2[keep]
        self.match_indices_1to2 = self.match_indices_1to2[keep]
        self.kp_colors = self.kp_colors[keep]

    def _legacy_estimate_camera_movement(self):
        _, self.R, self.t, _ = recover_pose(self.essential_matrix_1to2, self.points1_1to2,
                                            self.points2_1to2, self.image1.cameraMatrix)

    def _legacy_reconstruct_3d_points(self):
        K = self.image1.cameraMatrix
        self.null_projection_matrix = K @ np.eye(3, 4)
        # legacy quirk (:110): [R^T | t], not [R^T | -R^T t]
        self.projection_matrix = K @ np.hstack((self.R.T, self.t))
        X = triangulate_points(self.projection_matrix, self.null_projection_matrix,
                               self.points1_1to2.T, self.points2_1to2.T)
        X /= X[3, :]
        self.points3d_reconstr = X

    def get_validated_matches(self):
        self.filtered_matches_1to2 = [self.matches[i] for i in self.match_indices_1to2]
        kp1, kp2 = self.image1.keypoints, self.image2.keypoints
        self.points1_temp = [kp1[m.queryIdx].pt for m in self.filtered_matches_1to2]
        self.points2_temp = [kp2[m.trainIdx].pt for m in self.filtered_matches_1to2]
        self.points1_filtered = np.float32(self.points1_temp)
        self.points2_filtered = np.float32(self.points2_temp)

    def show_estimated_camera_motion(self):
        print("Estimated direction of translation between images")
        print(self.t)
        print("Estimated rotation of camera between images in degrees")
        print(rotationMatrixToEulerAngles(self.R) * 180 / math.pi)

    def visualize_filtered_matches(self):
        dm = [_cv2.DMatch(int(m.queryIdx), int(m.trainIdx), 0, float(m.distance))
              for m in self.filtered_matches_1to2]
        return _cv2.drawMatches(self.image1.image, self.image1.keypoints,
                                self.image2.image, self.image2.keypoints, dm, None,
                                flags=_cv2.DrawMatchesFlags_NOT_DRAW_SINGLE_POINTS)

    def visualise_points_in_3d_with_plotly(self):
        """Headless: the reference opens a plotly window (:133-211) — graphics are unavailable on
        this box; the point cloud stays in self.points3d_reconstr / self.kp_colors."""
        print("self.points3d_reconstr.shape")
        print(self.points3d_reconstr.shape)

    def standard_pipeline(self):
        self.match_detected_keypoints()
        self.determine_essential_matrix()
        self.estimate_camera_movement()
        self.get_validated_matches()
        self.reconstruct_3d_points()
