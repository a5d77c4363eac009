"""GPU: SiftDetector against the cv2 binary on the box (pyramid, key points, descriptors) + timing."""
import os
import sys
import time

import cv2
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import cv_sift as S  # noqa: E402  (checker only)
from visual_odometry_b200 import synth  # noqa: E402
from visual_odometry_b200.sift import SiftDetector  # noqa: E402


def cv_ref(img):
    kps, desc = cv2.SIFT_create().detectAndCompute(img, None)
    kp = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave) for k in kps],
                  np.float64).reshape(-1, 6)
    return kp, (np.zeros((0, 128), np.float32) if desc is None else desc)


def compare(kp, desc, ref, rdesc, tag):
    key = lambda a: {(round(float(r[0]), 2), round(float(r[1]), 2), round(float(r[2]), 2), round(float(r[3]), 1)): i
                     for i, r in enumerate(a)}
    a, b = key(kp), key(ref)
    common = sorted(set(a) & set(b))
    print(f"{tag}: ours {len(kp)} cv2 {len(ref)} common {len(common)} only_ours {len(set(a) - set(b))} "
          f"only_cv2 {len(set(b) - set(a))}", flush=True)
    if not common:
        return
    ia = np.array([a[k] for k in common]); ib = np.array([b[k] for k in common])
    print("   max |diff| x y size angle resp:", np.abs(kp[ia, :5] - ref[ib, :5]).max(0),
          "octave equal", np.array_equal(kp[ia, 5], ref[ib, 5]),
          "same order", np.array_equal(ia, ib) if len(kp) == len(ref) else None)
    dd = np.abs(desc[ia].astype(np.float32) - rdesc[ib])
    print("   desc: max", dd.max(), "values differing", float((dd > 0).mean()), "rows differing",
          float((dd.max(1) > 0).mean()), "rows > 2 levels", float((dd.max(1) > 2).mean()))


def main():
    det = SiftDetector()
    for (h, w, seed) in ((120, 160, 3), (480, 640, 11), (375, 1242, 7), (243, 325, 9)):
        img = synth.textured_image(h, w, seed)
        out = det.detect_batch(img[None], with_pyramid=True)
        torch.cuda.synchronize()
        n = int(out["n"][0]); st = int(out["status"][0])
        kp = out["kp"][0, :n].cpu().numpy().astype(np.float64); desc = out["desc"][0, :n].cpu().numpy()
        # pyramid: the oracle's (bit-exact against cv2.GaussianBlur, tests/test_oracle_sift.py)
        n_oct = S.n_octaves(h, w)
        gp = S.gaussian_pyramid(S.initial_image(img), n_oct)
        flat = np.concatenate([g.ravel() for g in gp])
        mine = out["pyramid"][0].cpu().numpy()
        print(f"{h}x{w}: status {st}, pyramid floats {mine.size} == {flat.size}, differing "
              f"{int((mine != flat).sum())}, max |diff| {np.abs(mine - flat).max()}", flush=True)
        ref, rdesc = cv_ref(img)
        compare(kp, desc, ref, rdesc, "vs cv2")
    # batch timing
    for (h, w, F) in ((480, 640, 32), (375, 1242, 32)):
        imgs = np.stack([synth.textured_image(h, w, 100 + i) for i in range(F)])
        d = torch.from_numpy(imgs).cuda()
        lib = det.detect_batch(d)                     # warm-up, allocates the workspace
        torch.cuda.synchronize()
        from visual_odometry_b200 import _lib
        L = _lib.load()
        L.vo_profile_enable(2)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        reps = 5
        for _ in range(reps):
            out = det.detect_batch(d)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        import ctypes as C
        parts = {}
        for kid, name in ((10, "pyramid"), (11, "extrema+orient+order"), (12, "describe")):
            tot, cnt = C.c_double(), C.c_int()
            L.vo_profile_read(kid, C.byref(tot), C.byref(cnt))
            parts[name] = round(tot.value / reps, 3)
        L.vo_profile_enable(0)
        nk = out["n"].cpu().numpy()
        t = time.time()
        for i in range(4):
            cv2.SIFT_create().detectAndCompute(imgs[i], None)
        cv_ms = (time.time() - t) / 4 * 1e3
        print(f"batch {F} x {h}x{w}: {ms:.3f} ms = {ms / F * 1e3:.1f} us/frame ({F / ms * 1e3:.0f} frames/s), "
              f"key points/frame {nk.mean():.0f}, parts ms {parts}; cv2 {cv_ms:.1f} ms/frame "
              f"({cv2.getNumThreads()} threads)", flush=True)


if __name__ == "__main__":
    main()
