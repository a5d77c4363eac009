"""Legacy two-image demo surface (reference src/image_and_keypoints.py:4-51 and
src/triangulate_points_from_images.py:7-41) on top of the B200 ImagePair."""
import numpy as np

from .cv2_shim import cv2
from .image_pair import ImagePair
from .matcher import B200Matcher, NORM_HAMMING, NORM_L2


class ImageAndKeypoints():
    def __init__(self, detector_name):
        orb = detector_name == "ORB"
        self.detector = cv2.ORB_create() if orb else cv2.SIFT_create()
        self.bf = B200Matcher(NORM_HAMMING if orb else NORM_L2, crossCheck=True)
        # the reference's effective intrinsics (../input/toys2/calibration.xml, :28-30);
        # the earlier DJI matrix at :19-26 is overwritten there too
        self.distCoeffs = np.array([[0.0097935857180804498, -0.021794052829051412,
                                     0.0046443590741258711, -0.0045664024579022498,
                                     0.017776502734846815]])
        self.cameraMatrix = np.array([[835.69, 0., 1008 / 2 + 61.6],
                                      [0., 827, 756 / 2 - 9.4],
                                      [0., 0., 1.]])
        self.scale_factor = 1
        self.cameraMatrix *= self.scale_factor
        self.cameraMatrix[2, 2] = 1

    def set_image(self, image):
        dim = (int(image.shape[1] * self.scale_factor), int(image.shape[0] * self.scale_factor))
        self.image = cv2.resize(image, dim, interpolation=cv2.INTER_AREA)

    def detect_keypoints(self):
        self.keypoints, self.descriptors = self.detector.detectAndCompute(self.image, None)
        self.kp_colors = [self.image[int(kp.pt[1]), int(kp.pt[0])] for kp in self.keypoints]


class TriangulatePointsFromTwoImages():
    def __init__(self):
        pass

    def load_images(self, filename_one, filename_two):
        self.img1 = cv2.imread(filename_one)
        self.img2 = cv2.imread(filename_two)

    def run(self, filename_one, filename_two, output="./output/images/filtered_matches.png"):
        self.load_images(filename_one, filename_two)
        detector = "ORB"
        views = []
        for img in (self.img1, self.img2):
            view = ImageAndKeypoints(detector)
            view.set_image(img)
            view.detect_keypoints()
            views.append(view)
        pair12 = ImagePair(detector)
        pair12.set_images(*views)
        pair12.standard_pipeline()
        pair12.show_estimated_camera_motion()
        pair12.visualise_points_in_3d_with_plotly()
        self.pair = pair12
        visualization12 = pair12.visualize_filtered_matches()
        cv2.imwrite(output, visualization12)
        cv2.waitKey(0)
        cv2.destroyAllWindows()
        return
