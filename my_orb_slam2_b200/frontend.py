"""Host-side mirror of the reference's front-end classes on top of the C ABI (include/orbfe.h).

Same names, argument meaning and error behaviour as the reference's C++ classes, so that the parity tests read
like calls into the reference:

  ORBextractor   include/ORBextractor.h:45-108   (operator() -> __call__, the six getters, mvImagePyramid)
  ORBmatcher     include/ORBmatcher.h:45-114     (DescriptorDistance, SearchByProjection x2, SearchForTriangulation)
  Frame          include/Frame.h (stereo ctor src/Frame.cc:72-130, ComputeStereoMatches :496-686)

The C++ drop-in classes for linking Tracking / LocalMapping / LoopClosing are in compat/; this Python mirror is the
harness used by tests/ and bench.py.  Everything is computed by liborbfe.so on the GPU; nothing here has a CPU path.
"""
import ctypes as C
import threading

import numpy as np

from . import orbfe
from .orbfe import KEYPOINT_DTYPE

TH_HIGH = 100        # src/ORBmatcher.cc:37
TH_LOW = 50          # :38
HISTO_LENGTH = 30    # :39
FRAME_GRID_COLS = 64  # include/Frame.h:37
FRAME_GRID_ROWS = 48  # :38


class ORBextractor:
    """ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) (src/ORBextractor.cc:410-413).

    The device handle is bound to an image size; it is (re)created on the first call / when the size changes."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=0, max_batch=1):
        self.nfeatures, self.scaleFactor, self.nlevels = nfeatures, scaleFactor, nlevels
        self.iniThFAST, self.minThFAST = iniThFAST, minThFAST
        self.device, self.max_batch = device, max_batch
        self._h = None
        self._pyr_cache = None
        # getters work before the first image: the tables do not depend on the image size
        probe = orbfe.Extractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, 64 * 4, 64 * 4, 1, device)
        self._info = probe.info
        probe.close()

    def handle(self, width, height, batch=1):
        if self._h is None or (self._h.width, self._h.height) != (width, height) or self._h.max_batch < batch:
            if self._h is not None:
                self._h.close()
            self._h = orbfe.Extractor(self.nfeatures, self.scaleFactor, self.nlevels, self.iniThFAST, self.minThFAST,
                                      width, height, max(batch, self.max_batch), self.device)
        return self._h

    def __call__(self, image, mask=None):
        """operator()(image, mask, keypoints, descriptors) (src/ORBextractor.cc:1065-1127); mask is ignored (:59)."""
        self._pyr_cache = None
        if image is None or image.size == 0:  # :1068 — silent return
            return np.zeros(0, KEYPOINT_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2  # :1072
        return self.handle(image.shape[1], image.shape[0]).extract(image)

    def extract_batch(self, images):
        """Many frames per call (not in the reference: its unit is one image)."""
        self._pyr_cache = None
        b, h, w = images.shape
        return self.handle(w, h, b).extract_batch(images)

    def GetLevels(self):
        return self.nlevels

    def GetScaleFactor(self):
        return self.scaleFactor

    def GetScaleFactors(self):
        return np.array(self._info.scale_factors[:self.nlevels], np.float32)

    def GetInverseScaleFactors(self):
        return np.array(self._info.inv_scale_factors[:self.nlevels], np.float32)

    def GetScaleSigmaSquares(self):
        return np.array(self._info.level_sigma2[:self.nlevels], np.float32)

    def GetInverseScaleSigmaSquares(self):
        return np.array(self._info.inv_level_sigma2[:self.nlevels], np.float32)

    @property
    def mnFeaturesPerLevel(self):
        return list(self._info.features_per_level[:self.nlevels])

    @property
    def mvImagePyramid(self):
        """Public member of the reference (include/ORBextractor.h:85); device resident, copied on demand."""
        if self._pyr_cache is None:
            if self._h is None:
                return []
            self._pyr_cache = [self._h.level(lv) for lv in range(self.nlevels)]
        return self._pyr_cache


class _View:
    def __init__(self, struct, arrays):
        self.struct = struct
        self.arrays = arrays


def _p(a):
    return a.ctypes.data_as(C.c_void_p).value


def frame_view(keys_un, u_right, desc, mp_obs, bounds, scale_factors):
    """The members of Frame that SearchByProjection reads (include/orbfe_types.h: orbfe_frame_view)."""
    a = [np.ascontiguousarray(keys_un, KEYPOINT_DTYPE), np.ascontiguousarray(u_right, np.float32),
         np.ascontiguousarray(desc, np.uint8), np.ascontiguousarray(mp_obs, np.int32),
         np.ascontiguousarray(scale_factors, np.float32)]
    min_x, min_y, max_x, max_y = [np.float32(v) for v in bounds]
    inv_w = np.float32(FRAME_GRID_COLS) / np.float32(max_x - min_x)  # src/Frame.cc:117-118
    inv_h = np.float32(FRAME_GRID_ROWS) / np.float32(max_y - min_y)
    s = orbfe.FrameView(len(a[0]), _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), min_x, min_y, max_x, max_y, inv_w, inv_h,
                        len(a[4]), _p(a[4]))
    return _View(s, a)


def mappoints_view(proj_x, proj_y, proj_xr, level, view_cos, in_view, obs, desc):
    a = [np.ascontiguousarray(proj_x, np.float32), np.ascontiguousarray(proj_y, np.float32),
         np.ascontiguousarray(proj_xr, np.float32), np.ascontiguousarray(level, np.int32),
         np.ascontiguousarray(view_cos, np.float32), np.ascontiguousarray(in_view, np.uint8),
         np.ascontiguousarray(obs, np.int32), np.ascontiguousarray(desc, np.uint8)]
    return _View(orbfe.MapPointsView(len(a[0]), *[_p(x) for x in a]), a)


def lastframe_view(valid, world_pos, obs, desc, octave, angle, tcw):
    a = [np.ascontiguousarray(valid, np.uint8), np.ascontiguousarray(world_pos, np.float32),
         np.ascontiguousarray(obs, np.int32), np.ascontiguousarray(desc, np.uint8),
         np.ascontiguousarray(octave, np.int32), np.ascontiguousarray(angle, np.float32)]
    t = (C.c_float * 12)(*[float(x) for x in np.asarray(tcw, np.float32).reshape(-1)[:12]])
    return _View(orbfe.LastFrameView(len(a[0]), *[_p(x) for x in a], t), a)


def camera_pose(tcw, fx, fy, cx, cy, mbf, mb):
    t = (C.c_float * 12)(*[float(x) for x in np.asarray(tcw, np.float32).reshape(-1)[:12]])
    return orbfe.CameraPose(t, fx, fy, cx, cy, mbf, mb)


def keyframe_view(keys_un, u_right, desc, has_mp, node_ids, node_offsets, node_features, scale_factors, level_sigma2):
    a = [np.ascontiguousarray(keys_un, KEYPOINT_DTYPE), np.ascontiguousarray(u_right, np.float32),
         np.ascontiguousarray(desc, np.uint8), np.ascontiguousarray(has_mp, np.uint8),
         np.ascontiguousarray(node_ids, np.uint32), np.ascontiguousarray(node_offsets, np.int32),
         np.ascontiguousarray(node_features, np.int32), np.ascontiguousarray(scale_factors, np.float32),
         np.ascontiguousarray(level_sigma2, np.float32)]
    s = orbfe.KeyFrameView(len(a[0]), _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), len(a[4]), _p(a[4]), _p(a[5]),
                           _p(a[6]), len(a[7]), _p(a[7]), _p(a[8]))
    return _View(s, a)


def points3d_view(valid, world_pos, normal, max_distance, min_distance, desc, angle=None):
    """Map points given by position (include/orbfe_types.h: orbfe_points3d_view).  max_distance / min_distance are
    MapPoint::mfMaxDistance / mfMinDistance; the two Get*DistanceInvariance() values are derived as MapPoint.cc:394-404
    does (1.2f * max, 0.8f * min in fp32)."""
    max_distance = np.ascontiguousarray(max_distance, np.float32)
    a = [np.ascontiguousarray(valid, np.uint8), np.ascontiguousarray(world_pos, np.float32),
         None if normal is None else np.ascontiguousarray(normal, np.float32),
         np.float32(1.2) * max_distance, np.float32(0.8) * np.ascontiguousarray(min_distance, np.float32), max_distance,
         np.ascontiguousarray(desc, np.uint8), None if angle is None else np.ascontiguousarray(angle, np.float32)]
    s = orbfe.Points3DView(len(a[0]), *[None if x is None else _p(x) for x in a])
    return _View(s, a)


def _f(vals, n):
    return (C.c_float * n)(*[float(x) for x in np.asarray(vals, np.float32).reshape(-1)[:n]])


def projection(tcw, fx, fy, cx, cy, bf, log_scale_factor, is_sim3=False, ow=None):
    """Pose (or Sim3) + calibration of the projection target (orbfe_projection)."""
    return orbfe.Projection(_f(tcw, 12), int(is_sim3), int(ow is not None), _f(ow if ow is not None else [0, 0, 0], 3),
                            fx, fy, cx, cy, bf, log_scale_factor)


def sim3_pair(t1w, t2w, s12, r12, t12, fx, fy, cx, cy, log_sf1, log_sf2):
    return orbfe.Sim3Pair(_f(t1w, 12), _f(t2w, 12), s12, _f(r12, 9), _f(t12, 3), fx, fy, cx, cy, log_sf1, log_sf2)


class ORBmatcher:
    """ORBmatcher(nnratio=0.6, checkOri=true) (include/ORBmatcher.h:45; src/ORBmatcher.cc:40-42)."""

    TH_LOW, TH_HIGH, HISTO_LENGTH = TH_LOW, TH_HIGH, HISTO_LENGTH
    # one C-ABI matcher handle (stream + workspace) per host thread and device: the reference's matcher objects are
    # stack-local and used concurrently by the Tracking, LocalMapping and LoopClosing threads; a
    # handle serves one call at a time (compat/ORBmatcher.cc keeps a thread_local handle the same way)
    _tls = threading.local()

    def __init__(self, nnratio=0.6, checkOri=True, device=0):
        self.mfNNratio, self.mbCheckOrientation, self.device = nnratio, checkOri, device

    @classmethod
    def _matcher(cls, device):
        ctx = cls._tls.__dict__.setdefault("ctx", {})
        if device not in ctx:
            ctx[device] = orbfe.Matcher(device)
        return ctx[device]

    @staticmethod
    def DescriptorDistance(a, b, device=0):
        """static int DescriptorDistance(const cv::Mat&, const cv::Mat&) (src/ORBmatcher.cc:1715-1731)."""
        return int(ORBmatcher._matcher(device).descriptor_distance(a, b)[0])

    def SearchByProjection(self, F, other, th=3.0, bMono=None, cam=None):
        """Two reference overloads share this name:
        (Frame&, const vector<MapPoint*>&, th) (src/ORBmatcher.cc:46)  -> other is a mappoints_view
        (Frame& Cur, const Frame& Last, th, bMono) (:1392)             -> other is a lastframe_view, cam required.
        Returns (assigned, nmatches): assigned[f] is the row written into F.mvpMapPoints[f] (-1 = untouched, -2 = reset to NULL by the
        rotation histogram of the last-frame overload)."""
        m = self._matcher(self.device)
        if isinstance(other.struct, orbfe.MapPointsView):
            return m.search_by_projection_map(F, other, th, self.mfNNratio)
        return m.search_by_projection_last(F, cam, other, th, bool(bMono), self.mbCheckOrientation)

    def SearchForTriangulation(self, kf1, kf2, F12, epipole, bOnlyStereo):
        """(KeyFrame*, KeyFrame*, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:702); returns the pair list."""
        return self._matcher(self.device).search_for_triangulation(kf1, kf2, F12, float(epipole[0]), float(epipole[1]),
                                                                   bOnlyStereo, self.mbCheckOrientation)

    def SearchByBoW(self, kf, other, kf_kf=False):
        """(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:182) when kf_kf is False -> matches[f] = index of
        the keyframe feature whose map point lands in vpMapPointMatches[f]; (KeyFrame*, KeyFrame*, vpMatches12) (:563)
        when True -> matches[idx1] = idx2.  Returns (matches, nmatches)."""
        return self._matcher(self.device).search_by_bow(kf, other, kf_kf, self.mfNNratio, self.mbCheckOrientation)

    def SearchForInitialization(self, F1, F2, vbPrevMatched, windowSize=10):
        """(Frame&, Frame&, vbPrevMatched, vnMatches12, windowSize) (:446); returns (vnMatches12, vbPrevMatched', n)."""
        return self._matcher(self.device).search_for_initialization(F1, F2, vbPrevMatched, windowSize, self.mfNNratio,
                                                                    self.mbCheckOrientation)

    def SearchByProjectionReloc(self, Cur, proj, pts, th, ORBdist):
        """(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) (:1540); returns (assigned, nmatches)."""
        return self._matcher(self.device).search_by_projection_reloc(Cur, proj, pts, th, ORBdist,
                                                                     self.mbCheckOrientation)

    def SearchByProjectionSim3(self, kf, proj, pts, th):
        """(KeyFrame*, Scw, vpPoints, vpMatched, th) (:321); returns (assigned, nmatches)."""
        return self._matcher(self.device).search_by_projection_sim3(kf, proj, pts, th)

    def SearchLocalPoints(self, F, proj, pts, obs, th=1.0, viewingCosLimit=0.5):
        """Tracking::SearchLocalPoints (src/Tracking.cc:1258-1284): Frame::isInFrustum over the local map fused with
        SearchByProjection(Frame&, vpMapPoints, th); returns (assigned, nmatches, per-point track fields)."""
        return self._matcher(self.device).search_local_points(F, proj, pts, obs, viewingCosLimit, th, self.mfNNratio)

    def Fuse(self, kf, proj, pts, th=3.0, inv_level_sigma2=None):
        """The search of both Fuse overloads (:879, :1033); returns (best_idx per point, nFused)."""
        return self._matcher(self.device).fuse(kf, proj, pts, th, inv_level_sigma2)

    def SearchBySim3(self, kf1, kf2, p1, p2, pair, th):
        """(KeyFrame*, KeyFrame*, vpMatches12, s12, R12, t12, th) (:1158); returns (match12, nFound)."""
        return self._matcher(self.device).search_by_sim3(kf1, kf2, p1, p2, pair, th)


class Frame:
    """Stereo Frame (src/Frame.cc:72-130): runs both extractors, then ComputeStereoMatches."""

    def __init__(self, imLeft, imRight, extractorLeft, extractorRight, bf, fx):
        self.mpORBextractorLeft, self.mpORBextractorRight = extractorLeft, extractorRight
        self.mbf = np.float32(bf)
        self.mb = np.float32(bf) / np.float32(fx)  # :127 (passed explicitly: SURVEY.md 8(c) deviation 2)
        self.mvKeys, self.mDescriptors = extractorLeft(imLeft)            # ExtractORB(0, imLeft)  (:89, :263)
        self.mvKeysRight, self.mDescriptorsRight = extractorRight(imRight)  # ExtractORB(1, imRight) (:90, :265)
        self.N = len(self.mvKeys)
        self.mvuRight = np.full(self.N, -1, np.float32)
        self.mvDepth = np.full(self.N, -1, np.float32)
        self.ComputeStereoMatches()

    def ComputeStereoMatches(self):
        """void Frame::ComputeStereoMatches() (src/Frame.cc:496-686) on the device-resident extractor outputs."""
        if self.N == 0:
            return 0
        ur, dp, n = orbfe.stereo_match(self.mpORBextractorLeft._h, self.mpORBextractorRight._h, 1, self.mbf, self.mb)
        self.mvuRight, self.mvDepth = ur[0][:self.N].copy(), dp[0][:self.N].copy()
        return int(n[0])


class MonoFrame:
    """Monocular Frame (src/Frame.cc:176-228): ExtractORB, UndistortKeyPoints, ComputeImageBounds.
    K = (fx, fy, cx, cy), distCoef = (k1, k2, p1, p2[, k3]) as Tracking builds them (src/Tracking.cc:58-100)."""

    def __init__(self, imGray, extractor, K, distCoef):
        self.mpORBextractorLeft = extractor
        self.cam = tuple(K) + tuple(distCoef)
        self.mvKeys, self.mDescriptors = extractor(imGray)                 # ExtractORB(0, imGray)  (:204)
        self.N = len(self.mvKeys)
        self.UndistortKeyPoints()                                          # :211
        self.mvuRight = np.full(self.N, -1, np.float32)                    # :214-215
        self.mvDepth = np.full(self.N, -1, np.float32)
        h, w = imGray.shape
        self.mnMinX, self.mnMaxX, self.mnMinY, self.mnMaxY = orbfe.image_bounds(self.cam, w, h)   # ComputeImageBounds (:222)

    def UndistortKeyPoints(self):
        """void Frame::UndistortKeyPoints() (src/Frame.cc:429-460) on the device-resident keypoints."""
        if self.N == 0:
            self.mvKeysUn = self.mvKeys.copy()
            return
        self.mvKeysUn = self.mpORBextractorLeft._h.undistort_keypoints(1, self.cam)[0][:self.N].copy()


class RGBDFrame(MonoFrame):
    """RGB-D Frame (src/Frame.cc:118-173): the mono steps + ComputeStereoFromRGBD(imDepth) (:688-712); imDepth is the
    CV_32F depth image after Tracking::GrabImageRGBD applied mDepthMapFactor."""

    def __init__(self, imGray, imDepth, extractor, K, distCoef, bf):
        super().__init__(imGray, extractor, K, distCoef)
        self.mbf = np.float32(bf)
        if self.N:
            ur, dz = self.mpORBextractorLeft._h.stereo_from_rgbd(1, imDepth, self.mbf)
            self.mvuRight, self.mvDepth = ur[0][:self.N].copy(), dz[0][:self.N].copy()
