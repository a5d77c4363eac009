"""The reference's SLAM frame loop on the device-resident path.

Mirrors VisualSlam.match_current_and_previous_frame (reference src/visual_slam.py:286-307) and what
it calls, frame by frame, on flat device arrays:
  hot path of every consecutive pair (match -> E-RANSAC -> recoverPose -> triangulation): ONE
  batched PairPipeline pass over the whole sequence (the pairs are independent);
  then per frame, in order: initialize_map (:43-87) for the first pair; update_feature_mapper,
  MatchWithMap -> solvePnPRansac (:190-235), add camera, add_information_to_map (:152-181: the
  pair's inliers re-triangulated with the two absolute camera poses), freeze_nonlast_cameras,
  optimize_map (40 LM iterations), remove observations with squared error >= 1 (:263-266),
  limit_number_of_camera_in_map(18) (:307).
Everything numerical runs in libvo_b200; the host only sequences launches and reads back a few
counters per frame.  Visualisation (imshow, ThreeDimViewer) is unavailable on this box.
"""
import numpy as np
import torch

from . import _lib
from .device_map import DeviceMap
from .pipeline import PairPipeline
from .pnp import CvError, rodrigues, solve_pnp_ransac_device

F64 = torch.float64


class DeviceSlam:
    def __init__(self, K, n_cap, max_frames, descriptor="sift", max_cameras=18, cap_points=None,
                 cap_obs=None, ba_iterations=40):
        self.K = np.asarray(K, np.float64).reshape(3, 3)
        self.n_cap, self.max_frames = int(n_cap), int(max_frames)
        self.max_cameras, self.ba_iterations = int(max_cameras), int(ba_iterations)
        self.pipe = PairPipeline(n_cap, max_frames - 1, max_frames, descriptor=descriptor, K=self.K)
        self.map = DeviceMap(self.K, max_frames, n_cap,
                             cap_points or max_frames * n_cap, cap_obs or 4 * max_frames * n_cap,
                             cap_cams=max_cameras + 2)
        self.lib = _lib.load()
        self.K_dev = torch.from_numpy(self.K).cuda()

    # ------------------------------------------------------------------ whole sequence
    def run(self, desc, kp, nkp):
        """desc [F,n_cap,D], kp [F,n_cap,2] float32, nkp [F] (host arrays) -> per-frame log."""
        F = len(nkp)
        pairs = np.stack([np.arange(F - 1), np.arange(1, F)], 1).astype(np.int32)
        self.pipe.run_host(desc, kp, nkp, pairs)
        return [self.step(b) for b in range(F - 1)]

    def run_images(self, images, detector):
        """The same loop fed with IMAGES (uint8 [F,H,W] or BGR [F,H,W,3]): feature extraction on
        the device (orb.OrbDetector or, for a "sift" loop, sift.SiftDetector; reference
        src/frame_generator.py:25) writes straight into the pipeline's frame arrays, then the hot
        path and the per-frame map stage as in `run`."""
        b = self.pipe.run_images(images, detector)
        return [self.step(k) for k in range(b)]

    # ------------------------------------------------------------------ streaming
    def add_frame(self, desc, kp):
        """VisualSlam.add_to_list_of_frames + match_current_and_previous_frame for ONE new frame
        (src/visual_slam.py:39-41, 286-307): upload its descriptors / keypoints, run the hot path
        on the pair (previous, new), then the map stage.  Returns None for the first frame."""
        p = self.pipe
        f = getattr(self, "_n_frames", 0)
        if f >= self.max_frames:
            # the track arrays are sized for max_frames * n_cap feature ids up front (flat device
            # arrays, no ring buffer): a stream longer than that needs a new DeviceSlam (the
            # reference's dicts grow without bound; its camera window is 18 frames)
            raise RuntimeError("DeviceSlam: max_frames exceeded (track arrays are sized up front)")
        n = int(desc.shape[0])
        p.desc[f, :n].copy_(torch.from_numpy(np.ascontiguousarray(desc)))
        p.kp[f, :n].copy_(torch.from_numpy(np.ascontiguousarray(kp, np.float32)))
        p.nkp[f] = n
        self._n_frames = f + 1
        if f == 0:
            return None
        p.pairs[f - 1].copy_(torch.tensor([f - 1, f], dtype=torch.int32))
        p.run_device(1, p0=f - 1, chain=False)
        return self.step(f - 1)

    # ------------------------------------------------------------------ one new frame (b + 1)
    def step(self, b):
        p, m = self.pipe, self.map
        q, t, mask, n = p.q[b], p.t_idx[b], p.mask[b], p.n_match[b:b + 1]
        info = dict(frame=b + 1, pnp_points=0, pnp_ok=None, ba=None)
        if b == 0:                                            # initialize_map
            m.initialize(q, t, mask, n, 0, 1, p.kp, p.X[0], p.R[0].cpu().numpy(), p.t[0].cpu().numpy())
            info["ba_init"] = m.optimize(self.ba_iterations)
        m.update_tracks(q, t, mask, n, b, b + 1)              # estimate_current_camera_position
        if b + 2 >= 3:
            try:
                img, xyz, _, _ = m.matches_with_map(q, t, mask, n, b, b + 1, p.kp)
                info["pnp_points"] = int(img.shape[0])
                ok, rvec, tvec, n_inl = solve_pnp_ransac_device(xyz, img, m.K)
                info["pnp_ok"], info["pnp_inliers"] = bool(ok), n_inl
                info["pnp_pose"] = np.r_[rvec.ravel(), tvec.ravel()]
                if ok:
                    m.add_camera(rodrigues(rvec)[0], tvec, b + 1, fixed=False)
                    self._add_information_to_map(b)
            except (CvError, ValueError, NotImplementedError) as e:   # "Position estimation failed"
                info["error"] = str(e)
            m.freeze_nonlast_cameras()
            info["ba"] = m.optimize(self.ba_iterations)
            m.prune(1.0)
        m.freeze_nonlast_cameras()
        m.limit_number_of_cameras(self.max_cameras)
        info["n_points"], info["n_obs"] = m.counts()
        info["cameras"] = list(m.cam_frame)
        return info

    def _add_information_to_map(self, b):
        """src/visual_slam.py:152-181: inliers of pair b re-triangulated with the absolute poses of
        the two cameras (cv2.triangulatePoints(K pose1, K pose2, pts1, pts2), /= w), then the map
        growth kernel with the 50-unit distance filter."""
        p, m = self.pipe, self.map
        c1, c2 = m.camera_row(b), m.camera_row(b + 1)         # ValueError = KeyError of camera_dict
        n = int(p.n_match[b].item())
        pts = p.pts[b, :n][p.mask[b, :n].bool()]
        k = int(pts.shape[0])
        X = torch.empty(max(k, 1), 4, dtype=F64, device=pts.device)
        if k:
            x1, x2 = pts[:, :2].contiguous(), pts[:, 2:].contiguous()
            Kt = self.K_dev
            P1 = (Kt @ m.poses[c1].view(4, 4)[:3]).contiguous()
            P2 = (Kt @ m.poses[c2].view(4, 4)[:3]).contiguous()
            _lib.check(self.lib.vo_triangulate(_lib.ptr(P1), _lib.ptr(P2), _lib.ptr(x1), _lib.ptr(x2),
                                               k, _lib.ptr(X), 1, _lib.cur_stream()))
        m.extend(p.q[b], p.t_idx[b], p.mask[b], p.n_match[b:b + 1], b, b + 1, p.kp, X)
