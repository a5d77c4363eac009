"""Frame record and FrameGenerator with the reference's call surface
(src/frame.py:4-14, src/frame_generator.py:5-38).

The detector (ORB/SIFT extraction) is outside the hot path: any object with
`detectAndCompute(image, None) -> (keypoints, descriptors)` works, including the synthetic
detector used by tests and bench (`SyntheticDetector`).  A packed copy of `KeyPoint.pt` is kept
on the frame so later stages never walk Python lists.
"""
import numpy as np

from .initials_types import Feature


class Frame():
    """Per-image record: image, id, keypoints, descriptors, features."""

    def __init__(self, image=None):
        self.image = image
        self.id = None
        self.keypoints = None
        self.descriptors = None
        self.features = None
        self.points = None  # float32 [N,2] packed KeyPoint.pt (extension; not in the reference)

    def __repr__(self):
        return repr('Frame %d' % (self.id))


class FrameGenerator():
    def __init__(self, detector):
        self.next_image_counter = 0
        self.detector = detector

    def make_frame(self, image) -> Frame:
        """Frame id = running counter; feature id = (frame id, keypoint index)
        (reference src/frame_generator.py:19-36)."""
        frame = Frame(image)
        frame.id = self.next_image_counter
        self.next_image_counter += 1
        frame.keypoints, frame.descriptors = self.detector.detectAndCompute(frame.image, None)
        frame.features = [Feature(kp, desc, (frame.id, k))
                          for k, (kp, desc) in enumerate(zip(frame.keypoints, frame.descriptors))]
        frame.points = np.array([kp.pt for kp in frame.keypoints], np.float32).reshape(-1, 2)
        return frame


class KeyPointLite:
    """Minimal cv2.KeyPoint stand-in (`.pt` tuple of two Python floats from float32)."""
    __slots__ = ("pt", "size", "angle", "response", "octave", "class_id")

    def __init__(self, x, y, size=31.0):
        self.pt = (float(np.float32(x)), float(np.float32(y)))
        self.size, self.angle, self.response, self.octave, self.class_id = size, -1.0, 0.0, 0, -1


class SyntheticDetector:
    """detectAndCompute over pre-generated (keypoints, descriptors): each call returns the next
    frame of a visual_odometry_b200.synth.Sequence.  Uses cv2.KeyPoint when cv2 is importable so
    the frames can be fed to cv2 drawing code too."""

    def __init__(self, sequence):
        self.seq = sequence
        self.cursor = 0
        try:
            import cv2
            self._kp = lambda x, y: cv2.KeyPoint(float(x), float(y), 31.0)
        except Exception:
            self._kp = KeyPointLite

    def detectAndCompute(self, image, mask=None):
        f = self.cursor
        self.cursor += 1
        n = int(self.seq.nkp[f])
        kps = tuple(self._kp(x, y) for x, y in self.seq.kp[f, :n])
        return kps, self.seq.desc[f, :n].copy()
