"""Headless counterpart of the reference's stream loop (`/root/reference/midas_mediapipe.py:11-160`): same class
name, constructor, method names and per-frame arithmetic, on a compute box without HighGUI / MediaPipe.

What is kept verbatim: the order of operations per frame (`:51-63`: detector.findIris -> estimator.predict ->
face.get_depth -> face.rel2abs), the face branch (`:64-95`: s2c distance from the iris diameter, head width from
the two HEAD mesh points) and the body branch (`:96-120`), `to_video_frame` (`:141-150`, wrapping uint8 `*5 +10`),
`side_by_side` (`:152-155`) and the performance dictionary keys.  What changes:
  * no windows, no `cv2.waitKey`, no `cv2.putText` overlays (text is returned in `self.messages` instead);
  * `src` is a video path (cv2.VideoCapture) or any iterable of uint8 BGR frames; `output` a path or None (frames
    are then kept in `self.outputs`);
  * per-stage timing uses `time.perf_counter` (wall clock).  The reference's `process_time` counts CPU time only and
    so reports a GPU-bound `predict` as almost free;
  * `stream(pipelined=True)` (default) keeps two frames in flight through `DepthEstimator.predict_stream`: the
    detector of frame i+1 runs on the host while the GPU works on frame i.  Frame order and every number handed
    to `FaceDet` are unchanged - `pipelined=False` is the reference's strictly serial loop, and the two are tested to
    agree bit for bit;
  * `stream(fused=True, batch=N)` additionally takes K8's fused outputs instead of the fp32 map: the two map values
    `FaceDet.get_depth` reads (face.py:106-107) are gathered on the device at the points `FaceDet.sample_indices`
    names, and `to_video_frame`'s wrapping uint8 grey level (:141-150) arrives as 1 byte per pixel - the full map never
    crosses PCIe and the host makes no pass over it.  N frames per submission amortise the per-launch latency of the
    small network.  The numbers handed to `FaceDet` are the same floats (K8's samples are bit-identical to the map).
The detector stays the reference's host code (`detectors.PersonDetector`, MediaPipe - not installable here); any
object with `findIris(frame)` / `findBody(frame)` / `visualize(frame)` / `HEAD` that fills `face.mesh`,
`face.l_iris`, `face.r_iris` works, which is what the tests use.
"""
from __future__ import annotations

from collections import deque
from statistics import median
from time import perf_counter

import numpy as np


def _resize_half(img: np.ndarray, size) -> np.ndarray:
    """cv2.resize(img, size) (bilinear default) when OpenCV is present, else plain 2x2 subsampling."""
    try:
        import cv2
        return cv2.resize(img, size)
    except ImportError:                                   # pragma: no cover - OpenCV is in the image
        w, h = size
        ys = (np.arange(h) * img.shape[0] / h).astype(int)
        xs = (np.arange(w) * img.shape[1] / w).astype(int)
        return img[ys][:, xs]


class _MapProxy:
    """What `FaceDet.get_depth` sees in place of the depth map when the stream runs fused: `.shape` and `[i, j]`.
    The two points requested from K8 come back with the batch; any other index (body-pose branch, indices numpy would
    wrap) is read from the device-resident map with numpy's indexing rules."""

    def __init__(self, shape, res, b, pts):
        self.shape, self._res, self._b, self._pts = tuple(shape), res, b, pts

    def __getitem__(self, ij):
        i, j = int(ij[0]), int(ij[1])
        if self._pts is not None:
            for k in (0, 1):
                if (i, j) == tuple(self._pts[k]):
                    return self._res.samples[self._b, k]
        H, W = self.shape
        if not (-H <= i < H and -W <= j < W):
            raise IndexError(f"index ({i}, {j}) is out of bounds for a map of shape {self.shape}")
        return np.float32(self._res.map_pixel(self._b, i % H, j % W))


class VidStream(object):
    '''
    a wrapper for OpenCV that accepts an estimater and a detector.
    '''

    def __init__(self, estimator, detector, face, src=None, output=None):
        self.detector = detector
        self.estimator = estimator
        self.face = face
        self.video = None
        self._frames = None
        self.w = self.h = 0
        if isinstance(src, str):
            import cv2
            self.video = cv2.VideoCapture(src)
            self.video.set(cv2.CAP_PROP_BUFFERSIZE, 2)
            self.w, self.h = int(self.video.get(3)), int(self.video.get(4))
        elif src is not None:
            self._frames = iter(src)
        self.FPS = 1 / 30
        self.FPS_MS = int(self.FPS * 1000)
        self.output = output
        self.writer = None
        self.outputs = []
        self.messages = []
        self.status = False
        self.frame = None
        self.cnt = 0
        self.performance = {'iris': list(), 'body': list(), 'depth': list()}
        self._det_state = None
        self.keep_outputs = True      # False: skip the side-by-side video frame (codec / overlay work is out of scope)

    # ------------------------------------------------------------------ frame source / sink
    def _read(self):
        if self.video is not None:
            if not self.video.isOpened():
                return False, None
            return self.video.read()
        try:
            f = next(self._frames)
        except (StopIteration, TypeError):
            return False, None
        if self.w == 0:
            self.h, self.w = f.shape[:2]
        return True, f

    def write_output(self, depth_frame):
        if not self.keep_outputs and self.output is None:
            return
        combo = self.side_by_side(self.frame, depth_frame)
        if self.output is not None:
            import cv2
            if self.writer is None:
                self.writer = cv2.VideoWriter(str(self.output), cv2.VideoWriter_fourcc(*'MJPG'), 20,
                                              (combo.shape[1], combo.shape[0]))
            self.writer.write(combo)
        else:
            self.outputs.append(combo)

    # ------------------------------------------------------------------ reference helpers (midas_mediapipe.py:131-160)
    def write_messages(self, messages, img):
        self.messages.append(list(messages))

    def to_video_frame(self, img):
        '''
        transforms midas depth frame to a video frame.
        '''
        with np.errstate(invalid="ignore"):
            output = img.astype(np.uint8)
        output *= 5
        output += 10
        return np.stack([output, output, output], axis=-1)          # cv2.merge([output, output, output])

    def side_by_side(self, img1, img2):
        frame1 = _resize_half(img1, (self.w // 2, self.h // 2))
        frame2 = _resize_half(img2, (self.w // 2, self.h // 2))
        return np.hstack((frame1, frame2))

    # ------------------------------------------------------------------ the per-frame tail (midas_mediapipe.py:61-120)
    def _consume(self, frame, mesh_state, depth_frame, vis=None):
        """Everything the reference does with a frame once its depth map exists.  `mesh_state` is the detector's
        result for THIS frame (captured when the detector ran), restored into `face` first so that pipelining does
        not mix the landmarks of two frames."""
        self.frame = frame
        self.face.mesh, self.face.l_iris, self.face.r_iris = mesh_state
        self.face.get_depth(depth_frame)
        self.face.rel2abs()
        if self.face.mesh is not None:
            if hasattr(self.detector, "visualize"):
                self.detector.visualize(self.frame)
            l_diameter = self.face.l_iris['radius'] * 2
            self.face.s2c_dist(self.face.w_iris, l_diameter)
            x1, y1 = self.face.mesh[self.detector.HEAD[0]]
            x2, y2 = self.face.mesh[self.detector.HEAD[1]]
            self.face.get_headw((x1, y1), (x2, y2))
            self.write_messages([f"S2C Distance (ft) - iris: {str(self.face.s2c_d)}",
                                 f'Head width (in): {str(round((self.face.head_w / 10) / 2.54, 2))}',
                                 f'head_w_mm: {str(self.face.head_w)}', f'focal length: {round(self.face.f, 2)}'],
                                self.frame)
        else:
            self.face.rel2abs()                                     # the reference calls it a second time here (:98)
            t0 = perf_counter()
            self.frame, head_pts = self.detector.findBody(self.frame)
            self.performance['body'].append(perf_counter() - t0)
            self.face.get_headw(head_pts[0], head_pts[1])
            self.face.s2c_dist(median(self.face.head_measurements), self.face.head_w)
        if vis is not None and not self.keep_outputs and self.output is None:
            depth_vis = vis                                         # consumer does not want the 3-channel video frame
        elif vis is not None:                                       # grey level already produced by K8 (same bytes)
            depth_vis = np.stack([vis, vis, vis], axis=-1)
        else:
            depth_vis = self.to_video_frame(depth_frame)
        self.write_messages([f'Relative Inverse Depth: {round(self.face.ri_depth, 2)}',
                             f'Absolute Depth: {round(self.face.abs_depth, 2)}',
                             f'RMSE: {round(self.face.rmse(), 2)}', f'MAE: {round(self.face.mae(), 2)}'], depth_vis)
        self.write_output(depth_vis)

    def _detect(self, frame):
        # the landmarks live in `face` (the detector writes them there) and persist across frames without a face;
        # when pipelined, `_consume` of an OLDER frame may have rewound them, so put the detector's latest view back
        if self._det_state is not None:
            self.face.l_iris, self.face.r_iris = dict(self._det_state[0]), dict(self._det_state[1])
        self.face.mesh = None
        t0 = perf_counter()
        self.detector.findIris(frame)
        self.performance['iris'].append(perf_counter() - t0)
        self._det_state = (dict(self.face.l_iris), dict(self.face.r_iris))
        return (self.face.mesh, dict(self.face.l_iris), dict(self.face.r_iris))

    # ------------------------------------------------------------------ the loop (midas_mediapipe.py:32-129)
    def stream(self, pipelined=True, max_frames=None, fused=False, batch=1):
        if fused and hasattr(self.estimator, "predict_stream_fused"):
            self._stream_fused(max_frames, max(1, int(batch)))
        elif pipelined and hasattr(self.estimator, "predict_stream"):
            self._stream_pipelined(max_frames)
        else:
            self._stream_serial(max_frames)
        if self.video is not None:
            self.video.release()
        if self.writer is not None:
            self.writer.release()
        return self.report()

    def _stream_serial(self, max_frames):
        while max_frames is None or self.cnt < max_frames:
            self.status, frame = self._read()
            if not self.status:
                break
            self.cnt += 1
            state = self._detect(frame)
            t0 = perf_counter()
            depth_frame = self.estimator.predict(frame)
            self.performance['depth'].append(perf_counter() - t0)
            self._consume(frame, state, depth_frame)

    def _stream_pipelined(self, max_frames):
        pending = deque()                                            # (frame, detector state, submit time)

        def feed():
            while max_frames is None or self.cnt < max_frames:
                self.status, frame = self._read()
                if not self.status:
                    return
                self.cnt += 1
                state = self._detect(frame)                          # host work of frame i+1 overlaps GPU work of frame i
                pending.append((frame, state, perf_counter()))
                yield frame[None]

        for maps in self.estimator.predict_stream(feed(), depth=2):
            frame, state, t0 = pending.popleft()
            self.performance['depth'].append(perf_counter() - t0)
            self._consume(frame, state, maps[0])

    def _stream_fused(self, max_frames, batch):
        pending = deque()                                            # (frames, detector states, requested points, t0)

        def points(state, shape):
            """(row, col) x 2 that `FaceDet.get_depth` will read for this frame, or None when they are not known at
            submission time (body-pose branch: the head points come from an earlier frame's findBody) or not valid
            numpy indices of the map (the reference would wrap / raise: left to `_MapProxy`)."""
            mesh, l_iris, r_iris = state
            if mesh is None:
                return None
            saved = (self.face.mesh, self.face.l_iris, self.face.r_iris)
            self.face.mesh, self.face.l_iris, self.face.r_iris = mesh, l_iris, r_iris
            try:
                a, b = self.face.sample_indices(shape)
            finally:
                self.face.mesh, self.face.l_iris, self.face.r_iris = saved
            ok = all(0 <= p[0] < shape[0] and 0 <= p[1] < shape[1] for p in (a, b))
            return (a, b) if ok else None

        def feed():
            while True:
                frames, states, pts = [], [], []
                while len(frames) < batch and (max_frames is None or self.cnt < max_frames):
                    self.status, frame = self._read()
                    if not self.status:
                        break
                    self.cnt += 1
                    state = self._detect(frame)
                    frames.append(frame)
                    states.append(state)
                    pts.append(points(state, frame.shape[:2]))
                if not frames:
                    return
                rc = np.zeros((len(frames), 2, 2), dtype=np.int32)
                for i, p in enumerate(pts):
                    if p is not None:
                        rc[i] = p
                pending.append((frames, states, pts, perf_counter()))
                yield frames, rc

        for res in self.estimator.predict_stream_fused(feed(), depth=2, n_samples=2, want_vis=True, want_map=False):
            frames, states, pts, t0 = pending.popleft()
            dt = (perf_counter() - t0) / len(frames)
            for b, (frame, state, p) in enumerate(zip(frames, states, pts)):
                self.performance['depth'].append(dt)
                self._consume(frame, state, _MapProxy(frame.shape[:2], res, b, p), vis=res.vis[b])

    def report(self):
        """The reference prints `1 // median(...)` FPS per stage at the end of the stream (:121-125)."""
        out = {}
        for k, v in self.performance.items():
            out[k] = (1.0 / median(v)) if v else None
        return out
