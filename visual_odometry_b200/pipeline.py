"""Batched two-view VO over many frame pairs: the form of the hot path that fills a B200.

One `PairPipeline.run_*` call = for every pair (a, b) of a frame set, exactly the reference's
per-pair chain (src/visual_slam.py:294-298 -> src/image_pair.py:233-354):
    match_features -> determine_essential_matrix -> estimate_camera_movement ->
    reconstruct_3d_points,
followed by the north-star trajectory chaining P_k = [R_k|t_k] P_{k-1} for consecutive pairs.
All stages are launches of libvo_b200 kernels on one CUDA stream; buffers are allocated once.
`run_host` is the end-to-end entry (host numpy in, host numpy out, copies inside).
"""
import ctypes
import os

import numpy as np
import torch

from . import _lib
from . import geometry as G

F64, I32, U8, F32 = torch.float64, torch.int32, torch.uint8, torch.float32


class _HostBuffer:
    """Owner of one vo_host_alloc block; the tensor made from it keeps it alive."""

    def __init__(self, nbytes, write_combined=True):
        self.lib = _lib.load()
        p = ctypes.c_void_p()
        _lib.check(self.lib.vo_host_alloc(ctypes.byref(p), nbytes, int(write_combined)))
        self.ptr, self.nbytes = p.value, nbytes

    def __del__(self):
        if getattr(self, "ptr", None):
            self.lib.vo_host_free(ctypes.c_void_p(self.ptr))
            self.ptr = None

    @staticmethod
    def tensor(owner, shape, dtype):
        n = int(np.prod(shape)) * torch.empty(0, dtype=dtype).element_size()
        hb = _HostBuffer(max(n, 1))
        raw = (ctypes.c_uint8 * hb.nbytes).from_address(hb.ptr)
        t = torch.frombuffer(raw, dtype=torch.uint8, count=n).view(dtype).reshape(*shape)
        owner.__dict__.setdefault("_host_blocks", []).append((hb, raw))
        return t


class PairPipeline:
    def __init__(self, n_cap, max_pairs, max_frames=None, descriptor="orb", K=None,
                 device="cuda", prob=0.99, threshold=1.0, max_iters=1000, adaptive=True,
                 match_mode=_lib.MATCH_MUTUAL, ratio=0.0, triangulate=True, chain=True):
        self.lib = _lib.load()
        self.dev = torch.device(device)
        self.n_cap, self.P = int(n_cap), int(max_pairs)
        self.F = int(max_frames if max_frames is not None else max_pairs + 1)
        self.descriptor = descriptor
        self.D = 32 if descriptor == "orb" else 128
        self.ddtype = U8 if descriptor == "orb" else F32
        self.prob, self.threshold = float(prob), float(threshold)
        self.max_iters, self.adaptive = int(max_iters), bool(adaptive)
        self.match_mode, self.ratio = int(match_mode), float(ratio)
        self.do_tri, self.do_chain = triangulate, chain
        dev, P, N, F = self.dev, self.P, self.n_cap, self.F
        e = lambda *s, dt: torch.empty(*s, dtype=dt, device=dev)
        # inputs resident in HBM; a second set (`slot` 1) is allocated on first use by the
        # streaming entry, so that the upload of batch k+1 overlaps the compute of batch k
        self._slots = [dict(desc=e(F, N, self.D, dt=self.ddtype), kp=e(F, N, 2, dt=F32),
                            nkp=e(F, dt=I32), pairs=e(P, 2, dt=I32))]
        self._use_slot(0)
        self.K = e(3, 3, dt=F64)
        if K is not None:
            self.K.copy_(torch.as_tensor(np.asarray(K, np.float64)))
        # per-pair outputs
        self.q, self.t_idx = e(P, N, dt=I32), e(P, N, dt=I32)
        self.dist = e(P, N, dt=F32)
        self.pts = e(P, N, 4, dt=F64)
        self.E = e(P, 3, 3, dt=F64)
        self.mask = e(P, N, dt=U8)
        self.n_iters = e(P, dt=I32)
        # what a caller reads back per batch (poses, counts) lies in ONE block so that the
        # streaming entry brings it to the host with one copy
        self._res_layout = self._result_layout(P)
        self._res = e(self._res_layout["bytes"], dt=U8)
        v = self._result_views(self._res, P)
        self.R, self.t, self.poses = v["R"], v["t"], v["poses"]
        self.n_inl, self.n_match, self.status = v["n_inl"], v["n_match"], v["status"]
        self.n_good = e(P, dt=I32)
        self.X = e(P, N, 4, dt=F64) if triangulate else None
        self.n_x = e(P, dt=I32)
        # workspaces
        if descriptor == "orb":
            self.ws_match = e(max(self.lib.vo_match_workspace_bytes(F, P, N), 256), dt=U8)
        else:
            self.ws_match = e(max(self.lib.vo_match_l2_workspace_bytes(F, P, N), 256), dt=U8)
        self.ws_ess = e(max(self.lib.vo_essential_workspace_bytes(P, N, self.max_iters), 256), dt=U8)
        self.ws_pose = e(max(self.lib.vo_pose_workspace_bytes(P, N), 256), dt=U8)
        # pinned staging for run_host
        self._pin = None
        self._copy_stream = None
        # kernels launched per run (for bench.py's gpu_launches claim; see launches_per_run)
        self.launches = 0

    @staticmethod
    def _result_layout(P):
        off, out = 0, {}
        for name, nbytes in (("R", P * 72), ("t", P * 24), ("poses", (P + 1) * 128),
                             ("n_inl", P * 4), ("n_match", P * 4), ("status", P * 4)):
            out[name] = (off, nbytes)
            off += (nbytes + 255) // 256 * 256      # sections stay 256-byte aligned like own allocations
        out["bytes"] = off
        return out

    def _result_views(self, raw, P):
        L = self._res_layout
        f = lambda n, dt, *shape: raw[L[n][0]:L[n][0] + L[n][1]].view(dt).view(*shape)
        return dict(R=f("R", F64, P, 3, 3), t=f("t", F64, P, 3), poses=f("poses", F64, P + 1, 4, 4),
                    n_inl=f("n_inl", I32, P), n_match=f("n_match", I32, P), status=f("status", I32, P))

    def _use_slot(self, k):
        while len(self._slots) <= k:
            a = self._slots[0]
            self._slots.append({n: torch.empty_like(v) for n, v in a.items()})
        sl = self._slots[k]
        self.desc, self.kp, self.nkp, self.pairs = sl["desc"], sl["kp"], sl["nkp"], sl["pairs"]
        self.slot = k

    # ------------------------------------------------------------------ device-resident run
    def run_device(self, n_pairs, stream=None, p0=0, chain=None, slot=None, after_match=None):
        """Inputs already in self.desc/kp/nkp/pairs. Asynchronous; results in the attributes.
        Processes pairs [p0, p0 + n_pairs) (outputs land in the same rows).  `after_match()` is
        called once the matching kernels are enqueued: the place to hang work on another stream
        that must not share the GPU with the persistent matching kernel (multi-GPU: the previous
        batch's all-gather — an NCCL kernel resident while the matching kernel starts keeps one
        of its 148 blocks waiting, and with it the kernel's end)."""
        if slot is not None and slot != self.slot:
            self._use_slot(slot)
        lib = self.lib
        st = _lib.cur_stream() if stream is None else stream
        N, B, a = self.n_cap, int(n_pairs), int(p0)
        p = lambda t: _lib.ptr(t[a:]) if a else _lib.ptr(t)      # per-pair arrays: row offset
        g = _lib.ptr                                             # frame-indexed / global arrays
        if B <= 0:
            return
        if self.descriptor == "orb":
            _lib.check(lib.vo_match_hamming_batch(
                g(self.desc), g(self.nkp), g(self.kp), N, self.F, p(self.pairs), B, self.match_mode,
                self.ratio, p(self.q), p(self.t_idx), p(self.dist), None, None, p(self.pts),
                p(self.n_match), g(self.ws_match), self.ws_match.numel(), st))
        else:
            _lib.check(lib.vo_match_l2_batch(
                g(self.desc), g(self.nkp), g(self.kp), N, self.D, self.F, p(self.pairs), B,
                self.match_mode, self.ratio, p(self.q), p(self.t_idx), p(self.dist), None, None,
                p(self.pts), p(self.n_match), g(self.ws_match), self.ws_match.numel(), st))
        if after_match is not None:
            after_match()
        _lib.check(lib.vo_essential_ransac_batch(
            p(self.pts), p(self.n_match), N, B, g(self.K), self.prob, self.threshold,
            self.max_iters, int(self.adaptive), p(self.E), p(self.mask), p(self.n_inl),
            p(self.n_iters), p(self.status), g(self.ws_ess), self.ws_ess.numel(), st))
        _lib.check(lib.vo_recover_pose_batch(
            p(self.E), p(self.pts), p(self.n_match), p(self.mask), N, B, g(self.K), 50.0,
            p(self.R), p(self.t), None, p(self.n_good), g(self.ws_pose), self.ws_pose.numel(), st))
        if self.do_tri:
            _lib.check(lib.vo_triangulate_pairs(
                p(self.R), p(self.t), p(self.pts), p(self.n_match), p(self.mask), N, B, g(self.K),
                p(self.X), p(self.n_x), st))
        if self.do_chain if chain is None else chain:
            _lib.check(lib.vo_chain_poses(g(self.R), g(self.t), a + B, None, g(self.poses), st))

    # ------------------------------------------------------------------ images in (N4 front end)
    def run_images(self, images, detector, pairs=None):
        """Frames as IMAGES: feature extraction (orb.OrbDetector.detect_batch, csrc/orb.cu, for an
        "orb" pipeline; sift.SiftDetector.detect_batch, csrc/sift.cu, for a "sift" one — the
        reference's live configuration, src/visual_slam.py:17-19) writes
        descriptors and key points straight into this pipeline's frame arrays on the device, then
        the hot path runs on `pairs` (default: consecutive frames).  What the reference does per
        frame in FrameGenerator.make_frame + match_current_and_previous_frame
        (src/frame_generator.py:10-38, src/visual_slam.py:286-307), for a whole sequence without a
        host round trip.  Returns the number of pairs; results are in the attributes."""
        out = detector.detect_batch(images, cap=self.n_cap)
        f = int(out["n"].shape[0])
        if f > self.F:
            raise ValueError("run_images: more frames than max_frames")
        if out["desc"].shape[-1] != self.D:
            raise ValueError("run_images: a %d-byte-descriptor detector on a '%s' pipeline"
                             % (out["desc"].shape[-1], self.descriptor))
        self.desc[:f].copy_(out["desc"])       # SIFT: uint8 -> the float32 rows cv2 would hand over
        self.kp[:f].copy_(out["kp"][:, :, :2])
        self.nkp[:f].copy_(out["n"])
        if pairs is None:
            pairs = np.stack([np.arange(f - 1), np.arange(1, f)], 1)
        pairs = torch.as_tensor(np.ascontiguousarray(pairs, np.int32))
        b = int(pairs.shape[0])
        self.pairs[:b].copy_(pairs)
        self.run_device(b)
        if int(out["status"].max().item()) != 0:
            raise _lib.VoError(_lib.VO_ERR_WORKSPACE,
                               "run_images: more than n_cap key points survived in some frame")
        return b

    def ransac_schedule(self, n_pairs=None):
        """iterations per RANSAC round, as vo_essential_ransac_batch schedules them"""
        import ctypes
        n_pairs = self.P if n_pairs is None else int(n_pairs)
        buf = (ctypes.c_int * 128)()
        n = self.lib.vo_ransac_round_schedule(self.max_iters, n_pairs, int(self.adaptive), buf, 128)
        return [int(buf[k]) for k in range(n)]

    def launches_per_run(self, n_pairs=None):
        """Number of libvo_b200 kernels one run_device issues (memsets excluded); mirrors the
        round schedule of vo_essential_ransac_batch (csrc/ransac.cu)."""
        n_pairs = self.P if n_pairs is None else int(n_pairs)
        rounds = len(self.ransac_schedule(n_pairs))
        if self.descriptor == "orb":
            # mirrors use_tensor_path (csrc/hamming.cu): expand + tcgen05 match + compact, or popc + compact
            ndir = 2 if self.match_mode == _lib.MATCH_MUTUAL else 1
            tc = self.n_cap >= 512 and n_pairs * ndir * ((self.n_cap + 255) // 256) >= 74
            match = 3 if tc else 2
        else:
            match = 5      # norms, tiles, GEMM (+ the launch that does not apply exits), rerank, compact
        return match + (1 + 4 * rounds + 1) + 1 + int(self.do_tri) + int(self.do_chain)

    # ------------------------------------------------------------------ host-buffer run (e2e)
    def _pinned(self):
        if self._pin is None:
            pin = lambda *s, dt: torch.empty(*s, dtype=dt, pin_memory=True)
            up = pin
            if os.environ.get("VO_B200_PINNED", "") == "wc":
                # upload-only buffers in write-combined page-locked memory (vo_host_alloc)
                up = lambda *s, dt: _HostBuffer.tensor(self, s, dt)
            self._pin = dict(
                desc=up(self.F, self.n_cap, self.D, dt=self.ddtype),
                kp=up(self.F, self.n_cap, 2, dt=F32), nkp=pin(self.F, dt=I32),
                pairs=pin(self.P, 2, dt=I32),
                R=pin(self.P, 3, 3, dt=F64), t=pin(self.P, 3, dt=F64),
                n_inl=pin(self.P, dt=I32), n_match=pin(self.P, dt=I32),
                status=pin(self.P, dt=I32), poses=pin(self.P + 1, 4, 4, dt=F64))
        return self._pin

    def stage_host(self, desc, kp, nkp, pairs):
        """Copy numpy inputs into the pinned staging area (not part of a timed step: it stands
        for the caller's own pinned frame buffers)."""
        pin = self._pinned()
        f, b = desc.shape[0], pairs.shape[0]
        pin["desc"][:f].copy_(torch.from_numpy(np.ascontiguousarray(desc)))
        pin["kp"][:f].copy_(torch.from_numpy(np.ascontiguousarray(kp, np.float32)))
        pin["nkp"][:f].copy_(torch.from_numpy(np.ascontiguousarray(nkp, np.int32)))
        pin["pairs"][:b].copy_(torch.from_numpy(np.ascontiguousarray(pairs, np.int32)))
        return f, b

    def run_staged(self, n_frames, n_pairs, chunks=2, first=0.15):
        """H2D of the staged inputs, the whole chain, D2H of poses/counts; synchronises.
        The frame set is uploaded in `chunks` pieces (the first one `first` of the frames) on a
        copy stream while the compute stream
        already works on the pairs whose frames have arrived (per-chunk CUDA events), so only the
        first piece of the upload is exposed.  Returns (h2d_bytes, d2h_bytes)."""
        pin, f, b = self._pinned(), int(n_frames), int(n_pairs)
        main = torch.cuda.current_stream()
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.dev)
        cs = self._copy_stream
        if self.descriptor != "orb":
            chunks = 1     # the L2 path converts the whole frame set to bf16 tiles per call
        chunks = max(1, min(int(chunks), f, max(b, 1)))
        # a small first piece (compute starts early), the rest in equal pieces
        f0 = max(1, round(first * f)) if chunks > 1 else f
        fb = [0] + [f0 + round(k * (f - f0) / (chunks - 1)) for k in range(chunks)] if chunks > 1 \
            else [0, f]
        cs.wait_stream(main)                       # earlier work may still read the buffers
        events = []
        with torch.cuda.stream(cs):
            self.nkp[:f].copy_(pin["nkp"][:f], non_blocking=True)
            self.pairs[:b].copy_(pin["pairs"][:b], non_blocking=True)
            for k in range(chunks):
                lo, hi = fb[k], fb[k + 1]
                self.desc[lo:hi].copy_(pin["desc"][lo:hi], non_blocking=True)
                self.kp[lo:hi].copy_(pin["kp"][lo:hi], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
                events.append(ev)
        # pair ranges: cut where the highest frame a pair needs crosses an upload boundary
        need = pin["pairs"][:b].numpy().max(axis=1) if b else np.zeros(0, np.int64)
        run_max = np.maximum.accumulate(need) if b else need
        p_lo = 0
        for k in range(chunks):
            p_hi = int(np.searchsorted(run_max, fb[k + 1], side="left")) if k + 1 < chunks else b
            main.wait_event(events[k])
            if p_hi > p_lo:
                self.run_device(p_hi - p_lo, p0=p_lo, chain=False)
                p_lo = p_hi
        if self.do_chain and b:
            _lib.check(self.lib.vo_chain_poses(_lib.ptr(self.R), _lib.ptr(self.t), b, None,
                                               _lib.ptr(self.poses), _lib.cur_stream()))
        pin["R"][:b].copy_(self.R[:b], non_blocking=True)
        pin["t"][:b].copy_(self.t[:b], non_blocking=True)
        pin["n_inl"][:b].copy_(self.n_inl[:b], non_blocking=True)
        pin["n_match"][:b].copy_(self.n_match[:b], non_blocking=True)
        pin["status"][:b].copy_(self.status[:b], non_blocking=True)
        pin["poses"][:b + 1].copy_(self.poses[:b + 1], non_blocking=True)
        main.synchronize()
        h2d = sum(pin[k][:f].numel() * pin[k].element_size() for k in ("desc", "kp", "nkp")) + \
            pin["pairs"][:b].numel() * 4
        d2h = b * (9 + 3) * 8 + b * 3 * 4 + (b + 1) * 16 * 8
        return h2d, d2h

    # ------------------------------------------------------------------ streamed batches (e2e)
    def run_stream(self, n_frames, n_pairs, steps, fill=None, consume=None, after_compute=None,
                   after_match=None, first0=0.15):
        """`steps` batches through the whole chain with the upload of batch k+1 running on a copy
        stream while batch k computes (two input slots on the device, two pinned result slots on
        the host).  Every batch: H2D of its frames from the pinned staging area, the chain, D2H of
        its poses / counts.  `fill(k, pin)` may refill the pinned input buffers for batch k (called
        once the previous upload has left them; default: the staged inputs are reused),
        `consume(k, res)` receives batch k's host results (views of a pinned slot, valid until
        batch k+2 is submitted), `after_compute(k)` may enqueue more work on the compute stream
        (multi-GPU: packing the poses), `after_match(k)` is passed to run_device (multi-GPU: the
        all-gather of batch k - 1).  Batch 0 has no predecessor to hide its upload behind: its frames
        go up in two pieces (the first `first0` of them) and the chain starts on the pairs whose
        frames have arrived.  Returns (h2d_bytes, d2h_bytes) per batch."""
        pin, f, b = self._pinned(), int(n_frames), int(n_pairs)
        main = torch.cuda.current_stream()
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.dev)
        cs = self._copy_stream
        if getattr(self, "_pin_out", None) is None:
            self._pin_out = []
            for _ in range(2):
                raw = torch.empty(self._res_layout["bytes"], dtype=U8, pin_memory=True)
                self._pin_out.append(dict(self._result_views(raw, self.P), raw=raw))
        self._use_slot(1)
        self._use_slot(0)
        uploaded = [None, None]       # event: inputs of the batch in this slot are on the device
        computed = [None, None]       # event: the chain of the batch in this slot has read them
        done = [None, None]           # event: results of the batch in this slot are on the host
        cs.wait_stream(main)

        def upload(k):
            sl = self._slots[k & 1]
            if fill is not None and not (k == 0 and filled0):
                if k > 0:
                    uploaded[(k - 1) & 1].synchronize()      # the staging area is free again
                fill(k, pin)
            with torch.cuda.stream(cs):
                if computed[k & 1] is not None:
                    cs.wait_event(computed[k & 1])           # batch k-2 no longer reads this slot
                sl["nkp"][:f].copy_(pin["nkp"][:f], non_blocking=True)
                sl["pairs"][:b].copy_(pin["pairs"][:b], non_blocking=True)
                sl["desc"][:f].copy_(pin["desc"][:f], non_blocking=True)
                sl["kp"][:f].copy_(pin["kp"][:f], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
            uploaded[k & 1] = ev

        def emit(k):
            done[k & 1].synchronize()
            if consume is not None:
                o = self._pin_out[k & 1]
                consume(k, dict(R=o["R"][:b], t=o["t"][:b], n_inl=o["n_inl"][:b],
                                n_match=o["n_match"][:b], status=o["status"][:b],
                                poses=o["poses"][:b + 1]))

        # Batch 0 has nothing to hide its upload behind: its frames go up in two pieces (the first
        # `first0` of them) and the chain starts on the pairs whose frames have arrived, as in
        # run_staged; every later batch is uploaded whole while its predecessor computes.
        first_cut = None
        filled0 = False
        if steps and self.descriptor == "orb" and f >= 64 and b >= 64:
            if fill is not None:
                fill(0, pin)               # the cut below reads batch 0's pair list
                filled0 = True
            f0 = max(2, round(first0 * f))
            need = np.maximum.accumulate(pin["pairs"][:b].numpy().max(axis=1))
            p0_hi = int(np.searchsorted(need, f0, side="left"))
            if 0 < p0_hi < b:
                first_cut = (f0, p0_hi)

        def upload_first():
            f0, _ = first_cut
            sl = self._slots[0]
            with torch.cuda.stream(cs):
                sl["nkp"][:f].copy_(pin["nkp"][:f], non_blocking=True)
                sl["pairs"][:b].copy_(pin["pairs"][:b], non_blocking=True)
                sl["desc"][:f0].copy_(pin["desc"][:f0], non_blocking=True)
                sl["kp"][:f0].copy_(pin["kp"][:f0], non_blocking=True)
                ev0 = torch.cuda.Event()
                ev0.record(cs)
                sl["desc"][f0:f].copy_(pin["desc"][f0:f], non_blocking=True)
                sl["kp"][f0:f].copy_(pin["kp"][f0:f], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
            uploaded[0] = ev
            return ev0

        ev_first = upload_first() if first_cut is not None else upload(0)
        for k in range(steps):
            if k + 1 < steps:
                upload(k + 1)
            am = (lambda kk=k: after_match(kk)) if after_match is not None else None
            if k == 0 and first_cut is not None:
                main.wait_event(ev_first)
                self.run_device(first_cut[1], slot=0, chain=False)
                main.wait_event(uploaded[0])
                self.run_device(b - first_cut[1], p0=first_cut[1], slot=0, chain=False, after_match=am)
                if self.do_chain:
                    _lib.check(self.lib.vo_chain_poses(_lib.ptr(self.R), _lib.ptr(self.t), b, None,
                                                       _lib.ptr(self.poses), _lib.cur_stream()))
            else:
                main.wait_event(uploaded[k & 1])
                self.run_device(b, slot=k & 1, after_match=am)
            ev = torch.cuda.Event()
            ev.record(main)
            computed[k & 1] = ev
            if after_compute is not None:
                after_compute(k)
            o = self._pin_out[k & 1]
            if b == self.P:
                o["raw"].copy_(self._res, non_blocking=True)        # one copy: poses and counts
            else:
                for name, src in (("R", self.R), ("t", self.t), ("n_inl", self.n_inl),
                                  ("n_match", self.n_match), ("status", self.status)):
                    o[name][:b].copy_(src[:b], non_blocking=True)
                o["poses"][:b + 1].copy_(self.poses[:b + 1], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(main)
            done[k & 1] = ev
            if k >= 1:
                emit(k - 1)
        if steps:
            emit(steps - 1)
        self._use_slot(0)
        h2d = sum(pin[n][:f].numel() * pin[n].element_size() for n in ("desc", "kp", "nkp")) + \
            pin["pairs"][:b].numel() * 4
        d2h = b * (9 + 3) * 8 + b * 3 * 4 + (b + 1) * 16 * 8
        if b == self.P:
            d2h = int(self._res_layout["bytes"])       # the block as copied (sections padded to 256 B)
        return h2d, d2h

    def run_host(self, desc, kp, nkp, pairs):
        """numpy in -> numpy out (R [B,3,3], t [B,3], n_inl, n_match, status, poses [B+1,4,4])."""
        f, b = self.stage_host(desc, kp, nkp, pairs)
        self.run_staged(f, b)
        pin = self._pin
        return dict(R=pin["R"][:b].numpy().copy(), t=pin["t"][:b].numpy().copy(),
                    n_inl=pin["n_inl"][:b].numpy().copy(),
                    n_match=pin["n_match"][:b].numpy().copy(),
                    status=pin["status"][:b].numpy().copy(),
                    poses=pin["poses"][:b + 1].numpy().copy())
