"""ctypes binding of the C ABI declared in include/ag2.h.

This is plumbing for tests/, bench.py and the multi-GPU host (sharded.py): it
marshals numpy arrays into the `ag2_*` entry points of libag2.so and nothing
else.  The same `Api` class binds the CPU oracle (prefix ``ag2o_``) for the
parity tests — the oracle itself is never loaded from this package
(see oracle/binding.py).

There is no CPU fallback: `load()` raises if libag2.so has not been built, and
`Context()` raises `Ag2Error(AG2_E_NO_DEVICE)` when no CUDA device is usable.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libag2.so")

AG2_OK = 0
AG2_E_INVALID, AG2_E_NO_DEVICE, AG2_E_CUDA, AG2_E_EMPTY_CLOUD = -1, -2, -3, -4
AG2_E_NO_SAMPLES, AG2_E_NO_WEIGHTS, AG2_E_STATE, AG2_E_IO, AG2_E_CAPACITY, AG2_E_NOMEM = -5, -6, -7, -8, -9, -10
MODE_NONE, MODE_PREDICTION, MODE_GEOMETRIC = 0, 1, 2
LIST_HYPOTHESES, LIST_PRUNED, LIST_ANTIPODAL, LIST_SELECTED = 0, 1, 2, 3
LIST_IMPORTANCE, LIST_CLUSTERS = 4, 5
IMAGE_BYTES = 60 * 60 * 3


class Params(C.Structure):
    """ag2_params (include/ag2.h) == HandSearch::Parameters + GraspDetector members."""

    _fields_ = [
        ("nn_radius_taubin", C.c_double),
        ("nn_radius_hands", C.c_double),
        ("num_orientations", C.c_int32),
        ("num_threads", C.c_int32),
        ("num_samples", C.c_int32),
        ("min_inliers", C.c_int32),
        ("finger_width", C.c_double),
        ("hand_outer_diameter", C.c_double),
        ("hand_depth", C.c_double),
        ("hand_height", C.c_double),
        ("init_bite", C.c_double),
        ("cam_tf_left", C.c_double * 16),
        ("cam_tf_right", C.c_double * 16),
        ("workspace", C.c_double * 6),
        ("antipodal_mode", C.c_int32),
        ("filter_half_grasps", C.c_int32),
        ("prune", C.c_int32),
        ("num_selected", C.c_int32),
        ("min_score_diff", C.c_double),
        ("min_aperture", C.c_double),
        ("max_aperture", C.c_double),
        ("normals_radius", C.c_double),
        ("rng_seed", C.c_uint64),
    ]

    def copy(self) -> "Params":
        p = Params()
        C.memmove(C.byref(p), C.byref(self), C.sizeof(Params))
        return p


HYPOTHESIS_DTYPE = np.dtype(
    [
        ("surface", "<f8", 3),
        ("bottom", "<f8", 3),
        ("top", "<f8", 3),
        ("axis", "<f8", 3),
        ("approach", "<f8", 3),
        ("binormal", "<f8", 3),
        ("width", "<f8"),
        ("score", "<f8"),
        ("sample_id", "<i4"),
        ("orientation_id", "<i4"),
        ("half_antipodal", "u1"),
        ("full_antipodal", "u1"),
        ("finger_slot", "u1"),
        ("depth_step", "u1"),
        ("num_in_box", "<i4"),
        ("hand_top", "<f8"),
        ("hand_bottom", "<f8"),
    ],
    align=True,
)
assert HYPOTHESIS_DTYPE.itemsize == 192

SWEEP_DTYPE = np.dtype(
    [
        ("fingers", "<u4"),
        ("hand", "<u2"),
        ("slot", "i1"),
        ("depth_step", "i1"),
        ("num_in_box", "<i4"),
        ("antipodal", "u1"),
        ("emitted", "u1"),
        ("reserved", "<u2"),
    ],
    align=True,
)
assert SWEEP_DTYPE.itemsize == 16

FRAME_DTYPE = np.dtype(
    [
        ("sample", "<f8", 3),
        ("normal", "<f8", 3),
        ("binormal", "<f8", 3),
        ("curvature_axis", "<f8", 3),
        ("valid", "<i4"),
        ("majority_cam", "<i4"),
        ("num_neighbors", "<i4"),
        ("reserved", "<i4"),
    ],
    align=True,
)
assert FRAME_DTYPE.itemsize == 112


class IsParams(C.Structure):
    """ag2_is_params (importance_sampling.cpp:9-15)."""

    _fields_ = [
        ("num_iterations", C.c_int32),
        ("num_samples", C.c_int32),
        ("num_init_samples", C.c_int32),
        ("method", C.c_int32),
        ("prob_rand_samples", C.c_double),
        ("radius", C.c_double),
        ("seed", C.c_uint64),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("num_points", C.c_int64),
        ("num_samples", C.c_int64),
        ("num_frames", C.c_int64),
        ("num_hypotheses", C.c_int64),
        ("num_pruned", C.c_int64),
        ("num_antipodal", C.c_int64),
        ("num_selected", C.c_int64),
        ("sum_neighbors_hands", C.c_int64),
        ("sum_cropped", C.c_int64),
        ("sum_in_box", C.c_int64),
        ("sum_neighbors_normals", C.c_int64),
        ("gpu_launches", C.c_int64),
        ("ms_phase", C.c_double * 8),
        ("ms_gap", C.c_double * 8),
    ]

    def as_dict(self) -> dict:
        d = {k: int(getattr(self, k)) for k, _ in self._fields_[:-2]}
        names = ["grid", "normals", "frames", "hands", "prune", "render", "lenet", "select"]
        d["ms_phase"] = dict(zip(names, list(self.ms_phase)))
        d["ms_gap"] = dict(zip(names, list(self.ms_gap)))
        return d


class Ag2Error(RuntimeError):
    def __init__(self, code: int, what: str, detail: str = ""):
        self.code = code
        super().__init__(f"{what} failed: {code}" + (f" ({detail})" if detail else ""))


LENET_SHAPES = [
    ("conv1_w", (20, 3, 5, 5)),
    ("conv1_b", (20,)),
    ("conv2_w", (50, 20, 5, 5)),
    ("conv2_b", (50,)),
    ("ip1_w", (500, 7200)),
    ("ip1_b", (500,)),
    ("ip2_w", (2, 500)),
    ("ip2_b", (2,)),
]


def _fptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Api:
    """Binds one shared library exporting the ag2.h surface under `prefix`."""

    def __init__(self, path: str, prefix: str = "ag2_"):
        if not os.path.exists(path):
            raise FileNotFoundError(
                f"{path} not found: build it first (python -c 'import __graft_entry__ as g; g.build()'). "
                "There is no CPU fallback for the CUDA path."
            )
        self.path = path
        self.prefix = prefix
        self.lib = C.CDLL(path, mode=C.RTLD_GLOBAL if prefix == "ag2_" else C.RTLD_LOCAL)

    def fn(self, name: str):
        f = getattr(self.lib, self.prefix + name)
        f.restype = C.c_int
        return f

    def has(self, name: str) -> bool:
        return hasattr(self.lib, self.prefix + name)


_api: Optional[Api] = None


def load() -> Api:
    """Loads libag2.so (the CUDA library).  Raises if it is missing."""
    global _api
    if _api is None:
        _preload_bundled_nccl()
        _api = Api(LIB_PATH, "ag2_")
    return _api


def _preload_bundled_nccl():
    """libag2.so needs libnccl.so.2 and so does torch, which ships a newer one than the system's: whichever is loaded
    first serves both.  Loading torch's copy first (when there is one) keeps a later `import torch` working in a
    process that loaded libag2.so before it; libag2 only uses calls that are the same in both."""
    import importlib.util
    try:
        spec = importlib.util.find_spec("nvidia.nccl")
    except Exception:
        spec = None
    for d in (spec.submodule_search_locations if spec else []):
        path = os.path.join(d, "lib", "libnccl.so.2")
        if os.path.exists(path):
            try:
                C.CDLL(path, mode=C.RTLD_GLOBAL)
            except OSError:
                pass
            return


def default_params() -> Params:
    p = Params()
    rc = load().fn("default_params")(C.byref(p))
    if rc:
        raise Ag2Error(rc, "ag2_default_params")
    return p


def read_lenet_caffemodel(path: str) -> dict:
    """ag2_read_lenet_caffemodel: the eight blobs of a .caffemodel as float32 arrays (host parser, no ctx)."""
    lib = load().lib
    out = {k: np.zeros(shp, np.float32) for k, shp in LENET_SHAPES}
    lib.ag2_read_lenet_caffemodel.restype = C.c_int
    rc = lib.ag2_read_lenet_caffemodel(path.encode(), *[_fptr(out[k]) for k, _ in LENET_SHAPES])
    if rc:
        raise Ag2Error(rc, "ag2_read_lenet_caffemodel", path)
    return out


def pcd_read(path: str):
    """ag2_pcd_read: returns (xyz [N,3] f32, normals [N,3] f32 or None, rgba [N] u32 or None)."""
    lib = load().lib
    n, hn, hc = C.c_size_t(), C.c_int(), C.c_int()
    rc = lib.ag2_pcd_info(path.encode(), C.byref(n), C.byref(hn), C.byref(hc))
    if rc:
        raise Ag2Error(rc, "ag2_pcd_info", path)
    N = int(n.value)
    xyz = np.zeros((max(N, 1), 3), np.float32)
    nrm = np.zeros((max(N, 1), 3), np.float32) if hn.value else None
    col = np.zeros(max(N, 1), np.uint32) if hc.value else None
    rc = lib.ag2_pcd_read(path.encode(), _fptr(xyz), _fptr(nrm), _fptr(col), C.c_size_t(max(N, 1)), C.byref(n))
    if rc:
        raise Ag2Error(rc, "ag2_pcd_read", path)
    return xyz[:N], (None if nrm is None else nrm[:N]), (None if col is None else col[:N])


def pcd_write(path: str, xyz, normals=None, rgba=None, binary=True):
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    nrm = None if normals is None else np.ascontiguousarray(normals, np.float32)
    col = None if rgba is None else np.ascontiguousarray(rgba, np.uint32)
    rc = load().lib.ag2_pcd_write(path.encode(), _fptr(xyz), _fptr(nrm), _fptr(col), C.c_size_t(len(xyz)),
                                  C.c_int(int(binary)))
    if rc:
        raise Ag2Error(rc, "ag2_pcd_write", path)


def merge_selected_blocks(blocks: np.ndarray, num_selected: int) -> np.ndarray:
    """ag2_merge_selected: blocks [world, cap] records of an all-gather -> the single-GPU selection."""
    blocks = np.ascontiguousarray(blocks, dtype=HYPOTHESIS_DTYPE)
    world, cap = blocks.shape
    out = np.zeros(world * cap, dtype=HYPOTHESIS_DTYPE)
    n = C.c_size_t()
    rc = load().fn("merge_selected")(blocks.ctypes.data_as(C.c_void_p), C.c_int(world), C.c_size_t(cap),
                                     C.c_int(int(num_selected)), out.ctypes.data_as(C.c_void_p), C.byref(n))
    if rc:
        raise Ag2Error(rc, "merge_selected")
    return out[: int(n.value)]


def serialize_grasp_list(hands: np.ndarray, seq=0, sec=0, nsec=0, frame_id="") -> bytes:
    hands = np.ascontiguousarray(hands, dtype=HYPOTHESIS_DTYPE)
    ln = C.c_size_t()
    lib = load().lib
    lib.ag2_serialize_grasp_list(_fptr(hands), C.c_size_t(len(hands)), C.c_uint32(seq), C.c_uint32(sec),
                                 C.c_uint32(nsec), frame_id.encode(), None, C.c_size_t(0), C.byref(ln))
    out = np.zeros(max(int(ln.value), 1), np.uint8)
    rc = lib.ag2_serialize_grasp_list(_fptr(hands), C.c_size_t(len(hands)), C.c_uint32(seq), C.c_uint32(sec),
                                      C.c_uint32(nsec), frame_id.encode(), _fptr(out), C.c_size_t(out.size),
                                      C.byref(ln))
    if rc:
        raise Ag2Error(rc, "ag2_serialize_grasp_list")
    return out[: int(ln.value)].tobytes()


class Context:
    """One ag2_ctx (or ag2o_ctx): one GPU, one in-flight detect call."""

    def __init__(self, params: Params, device: int = 0, api: Optional[Api] = None):
        self.api = api or load()
        self.is_oracle = self.api.prefix != "ag2_"
        self._h = C.c_void_p()
        self.params = params.copy()
        if self.is_oracle:
            rc = self.api.fn("create")(C.byref(self.params), C.byref(self._h))
        else:
            rc = self.api.fn("create")(C.byref(self.params), C.c_int(device), C.byref(self._h))
        if rc:
            raise Ag2Error(rc, self.api.prefix + "create")
        self.n_points = 0
        self.n_samples = 0

    # -- plumbing
    def _call(self, name: str, *args):
        rc = self.api.fn(name)(self._h, *args)
        if rc:
            detail = ""
            if not self.is_oracle and self.api.has("last_error"):
                f = self.api.lib.ag2_last_error
                f.restype = C.c_char_p
                detail = (f(self._h) or b"").decode()
            raise Ag2Error(rc, self.api.prefix + name, detail)

    def close(self):
        if self._h:
            self.api.fn("destroy")(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- parameters / weights
    def set_params(self, params: Params):
        self.params = params.copy()
        self._call("set_params", C.byref(self.params))

    def set_lenet_weights(self, weights: dict):
        arrs = []
        for name, shape in LENET_SHAPES:
            a = np.ascontiguousarray(weights[name], dtype=np.float32)
            assert a.shape == shape, (name, a.shape, shape)
            arrs.append(a)
        self._keep = arrs
        self._call("set_lenet_weights", *[_fptr(a) for a in arrs])

    def init_lenet_xavier(self, seed: int):
        self._call("init_lenet_xavier", C.c_uint64(seed))

    def load_lenet_caffemodel(self, path: str):
        self._call("load_lenet_caffemodel", C.c_char_p(path.encode()))

    def get_lenet_weights(self) -> dict:
        out = {n: np.zeros(s, dtype=np.float32) for n, s in LENET_SHAPES}
        self._call("get_lenet_weights", *[_fptr(out[n]) for n, _ in LENET_SHAPES])
        return out

    # -- inputs
    def set_cloud(self, xyz: np.ndarray, cam_mask: Optional[np.ndarray] = None, num_cams: int = 1,
                  normals: Optional[np.ndarray] = None):
        xyz = np.asarray(xyz)
        assert xyz.dtype == np.float32 and xyz.ndim == 2 and xyz.shape[1] >= 3
        if not xyz.flags.c_contiguous:
            xyz = np.ascontiguousarray(xyz)
        n = xyz.shape[0]
        stride = xyz.strides[0]
        cam = None if cam_mask is None else np.ascontiguousarray(cam_mask, dtype=np.uint8)
        nrm = None if normals is None else np.ascontiguousarray(normals, dtype=np.float32)
        if cam is not None:
            assert cam.shape == (n,)
        if nrm is not None:
            assert nrm.shape == (n, 3)
        self._cloud_keep = (xyz, cam, nrm)
        self._call("set_cloud", _fptr(xyz), C.c_size_t(stride), C.c_size_t(n), _fptr(cam), C.c_int(num_cams),
                   _fptr(nrm))
        self.n_points = n

    def set_cloud_device(self, d_xyz: int, n: int, d_cam_mask: int, num_cams: int, d_normals: int = 0):
        self._call("set_cloud_device", C.c_void_p(d_xyz), C.c_size_t(n), C.c_void_p(d_cam_mask or None),
                   C.c_int(num_cams), C.c_void_p(d_normals or None))
        self.n_points = n

    def set_sample_indices(self, idx):
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        self._call("set_sample_indices", _fptr(idx), C.c_size_t(idx.size))
        self.n_samples = idx.size

    def set_sample_indices_device(self, d_ptr: int, n: int):
        """Sample indices already in device memory (int32), e.g. after an NCCL broadcast."""
        self._call("set_sample_indices_device", C.c_void_p(d_ptr), C.c_size_t(n))
        self.n_samples = n

    def set_sample_points(self, xyz):
        xyz = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
        self._call("set_sample_points", _fptr(xyz), C.c_size_t(xyz.shape[0]))
        self.n_samples = xyz.shape[0]

    def set_sample_id_offset(self, first: int):
        self._call("set_sample_id_offset", C.c_int64(first))

    def set_sample_id_map(self, first: int, stride: int):
        """Sample s of this context is sample first + stride*s of the whole list (strided shard)."""
        self._call("set_sample_id_map", C.c_int64(first), C.c_int64(stride))

    # -- stages
    def compute_normals(self):
        self._call("compute_normals")

    def generate_hypotheses(self):
        self._call("generate_hypotheses")

    def prune(self):
        self._call("prune")

    def render_images(self, fetch: bool = True) -> Optional[np.ndarray]:
        if not fetch:
            self._call("render_images", None)
            return None
        # the count is known after prune
        n = self.num(LIST_PRUNED)
        out = np.zeros((n, 60, 60, 3), dtype=np.uint8)
        self._call("render_images", _fptr(out))
        return out

    def classify(self):
        self._call("classify")

    def select(self):
        self._call("select")

    def detect(self):
        self._call("detect")

    def preprocess_cloud(self, xyz: np.ndarray, cam_mask: Optional[np.ndarray] = None, num_cams: int = 1,
                         normals: Optional[np.ndarray] = None, filter_workspace: bool = True,
                         voxelize: bool = True, voxel_size: float = 0.003):
        """filterWorkspace + voxelizeCloud on the original cloud; the result becomes the ctx's cloud."""
        xyz = np.ascontiguousarray(xyz, dtype=np.float32).reshape(-1, 3)
        cam = None if cam_mask is None else np.ascontiguousarray(cam_mask, dtype=np.uint8)
        nrm = None if normals is None else np.ascontiguousarray(normals, dtype=np.float32)
        self._call("preprocess_cloud", _fptr(xyz), C.c_size_t(12), C.c_size_t(len(xyz)), _fptr(cam), C.c_int(num_cams),
                   _fptr(nrm), C.c_int(int(filter_workspace)), C.c_int(int(voxelize)), C.c_double(voxel_size))
        self.n_points = len(self.cloud()[0])

    def cloud(self):
        n = C.c_size_t()
        self._call("get_cloud", None, None, C.c_size_t(1 << 40), C.byref(n))
        xyz = np.zeros((int(n.value), 3), dtype=np.float32)
        cam = np.zeros(int(n.value), dtype=np.uint8)
        if n.value:
            self._call("get_cloud", _fptr(xyz), _fptr(cam), C.c_size_t(int(n.value)), C.byref(n))
        return xyz, cam

    def subsample_uniformly(self, num_samples: int, seed: int = 0) -> np.ndarray:
        self._call("subsample_uniformly", C.c_int(num_samples), C.c_uint64(seed))
        n = C.c_size_t()
        idx = np.zeros(max(num_samples, 1), dtype=np.int32)
        self._call("get_sample_indices", _fptr(idx), C.c_size_t(idx.size), C.byref(n))
        self.n_samples = int(n.value)
        return idx[: int(n.value)].copy()

    def sample_indices(self) -> np.ndarray:
        n = C.c_size_t()
        idx = np.zeros(max(self.n_samples, 1), dtype=np.int32)
        self._call("get_sample_indices", _fptr(idx), C.c_size_t(idx.size), C.byref(n))
        return idx[: int(n.value)].copy()

    def render_point_sets(self, pts, normals, offsets, aligns: bool = False) -> np.ndarray:
        """Learning::createGraspImage(points, normals, grasp): images of explicit unit-box point sets."""
        pts = np.ascontiguousarray(pts, dtype=np.float64).reshape(-1, 3)
        normals = np.ascontiguousarray(normals, dtype=np.float64).reshape(-1, 3)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
        n = len(offsets) - 1
        out = np.zeros((max(n, 1), 60, 60, 3), dtype=np.uint8)
        self._call("render_point_sets", _fptr(pts), _fptr(normals), _fptr(offsets), C.c_size_t(n),
                   C.c_int(1 if aligns else 0), _fptr(out))
        return out[:n]

    def set_min_inliers(self, min_inliers: int):
        """HandleSearch::setMinInliers: the cluster threshold alone, without invalidating the lists."""
        self._call("set_min_inliers", C.c_int(min_inliers))

    def find_clusters(self, which_src: int):
        """HandleSearch::findClusters on list `which_src`; result in LIST_CLUSTERS."""
        self._call("find_clusters", C.c_int(which_src))

    def default_is_params(self) -> IsParams:
        ip = IsParams()
        rc = self.api.fn("default_is_params")(C.byref(ip))
        if rc:
            raise Ag2Error(rc, "default_is_params")
        return ip

    def detect_importance(self, ip: IsParams):
        """ImportanceSampling::detectGraspPoses on the samples currently set; result in LIST_IMPORTANCE."""
        self._call("detect_importance", C.byref(ip))
        self.n_samples = int(ip.num_samples)

    def draw_importance_samples(self, ip: IsParams, rnd: int, hands: np.ndarray) -> np.ndarray:
        """The sample points of round `rnd` drawn from `hands` (all hands found so far), as
        detect_importance draws them; [num_samples, 3] float64."""
        hands = np.ascontiguousarray(hands, dtype=HYPOTHESIS_DTYPE)
        out = np.zeros((int(ip.num_samples), 3), dtype=np.float64)
        self._call("draw_importance_samples", C.byref(ip), C.c_int(rnd), hands.ctypes.data_as(C.c_void_p),
                   C.c_size_t(len(hands)), _fptr(out))
        return out

    def importance_samples(self, rnd: int) -> np.ndarray:
        n = C.c_size_t()
        out = np.zeros((1 << 20, 3), dtype=np.float64)
        self._call("get_importance_samples", C.c_int(rnd), _fptr(out), C.c_size_t(out.shape[0]), C.byref(n))
        return out[: int(n.value)].copy()

    def synchronize(self):
        if not self.is_oracle:
            self._call("synchronize")

    def measure_fp_peaks(self):
        """(fp32, fp64) FMA throughput of this ctx's GPU in TFLOP/s (ag2_measure_fp_peaks)."""
        a, b = C.c_double(), C.c_double()
        self._call("measure_fp_peaks", C.byref(a), C.byref(b))
        return float(a.value), float(b.value)

    def stream(self) -> int:
        h = C.c_void_p()
        self._call("get_stream", C.byref(h))
        return int(h.value or 0)

    # -- outputs
    def num(self, which: int) -> int:
        n = C.c_size_t()
        self._call("num_hypotheses", C.c_int(which), C.byref(n))
        return int(n.value)

    def hypotheses(self, which: int = LIST_HYPOTHESES) -> np.ndarray:
        n = self.num(which)
        out = np.zeros(max(n, 1), dtype=HYPOTHESIS_DTYPE)
        got = C.c_size_t()
        self._call("get_hypotheses", C.c_int(which), _fptr(out), C.c_size_t(out.size), C.byref(got))
        return out[: int(got.value)]

    def normals_shard_points(self, world: int) -> int:
        n = C.c_size_t()
        self._call("normals_shard_points", C.c_int(world), C.byref(n))
        return int(n.value)

    def compute_normals_shard(self, rank: int, world: int, d_out: int, cap_points: int) -> int:
        n = C.c_size_t()
        self._call("compute_normals_shard", C.c_int(rank), C.c_int(world), C.c_void_p(d_out), C.c_size_t(cap_points),
                   C.byref(n))
        return int(n.value)

    def set_normals_shards(self, d_all: int, cap_points: int, world: int):
        self._call("set_normals_shards", C.c_void_p(d_all), C.c_size_t(cap_points), C.c_int(world))

    def copy_hypotheses_device(self, which: int, d_ptr: int, cap: int) -> int:
        """Copies list `which` into caller-owned device memory of `cap` records (rest zero-filled)."""
        got = C.c_size_t()
        self._call("copy_hypotheses_device", C.c_int(which), C.c_void_p(d_ptr), C.c_size_t(cap), C.byref(got))
        return int(got.value)

    def stats(self) -> dict:
        s = Stats()
        self._call("get_stats", C.byref(s))
        return s.as_dict()

    def normals(self) -> np.ndarray:
        out = np.zeros((self.n_points, 3), dtype=np.float32)
        self._call("get_normals", _fptr(out))
        return out

    def frames(self) -> np.ndarray:
        out = np.zeros(max(self.n_samples, 1), dtype=FRAME_DTYPE)
        self._call("get_frames", _fptr(out), C.c_size_t(out.size))
        return out[: self.n_samples]

    def sweep_records(self) -> np.ndarray:
        R = self.params.num_orientations
        out = np.zeros(max(self.n_samples * R, 1), dtype=SWEEP_DTYPE)
        self._call("get_sweep_records", _fptr(out), C.c_size_t(out.size))
        return out[: self.n_samples * R].reshape(self.n_samples, R)

    def radius_search(self, q, radius: float, brute: bool = False):
        q = np.ascontiguousarray(q, dtype=np.float32)
        cap = max(self.n_points, 1)
        idx = np.zeros(cap, dtype=np.int32)
        d2 = np.zeros(cap, dtype=np.float32)
        n = C.c_size_t()
        if self.is_oracle:
            self._call("radius_search", _fptr(q), C.c_double(radius), C.c_int(int(brute)), _fptr(idx), _fptr(d2),
                       C.c_size_t(cap), C.byref(n))
        else:
            self._call("radius_search", _fptr(q), C.c_double(radius), _fptr(idx), _fptr(d2), C.c_size_t(cap),
                       C.byref(n))
        k = int(n.value)
        return idx[:k].copy(), d2[:k].copy()

    def in_box(self, which: int, h: int) -> np.ndarray:
        cap = max(self.n_points, 1)
        idx = np.zeros(cap, dtype=np.int32)
        n = C.c_size_t()
        self._call("get_in_box", C.c_int(which), C.c_size_t(h), _fptr(idx), C.c_size_t(cap), C.byref(n))
        return idx[: int(n.value)].copy()

    def logits(self) -> np.ndarray:
        n = self.num(LIST_PRUNED)
        out = np.zeros((max(n, 1), 2), dtype=np.float32)
        self._call("get_logits", _fptr(out), C.c_size_t(out.shape[0]))
        return out[:n]

    def classify_images(self, images: np.ndarray) -> np.ndarray:
        images = np.ascontiguousarray(images, dtype=np.uint8).reshape(-1, 60, 60, 3)
        out = np.zeros((images.shape[0], 2), dtype=np.float32)
        self._call("classify_images", _fptr(images), C.c_size_t(images.shape[0]), _fptr(out))
        return out

    def classify_images_static_ptr(self, d_images: int, n: int, d_logits: int) -> None:
        """ag2_classify_images_static: device pointers in and out, fixed scratch inside libag2.so, no
        allocation; enqueued on the ctx stream (synchronize() before reading the logits)."""
        self._call("classify_images_static", C.c_void_p(d_images), C.c_size_t(n), C.c_void_p(d_logits))


class _BorrowedContext(Context):
    """An ag2_ctx owned by a Group (weights, statistics, stream); closing it does nothing."""

    def __init__(self, handle, params: Params, api: Api):
        self.api = api
        self.is_oracle = False
        self._h = handle
        self.params = params.copy()
        self.n_points = 0
        self.n_samples = 0

    def close(self):
        self._h = C.c_void_p()


class GroupId(C.Structure):
    """ag2_group_id: an ncclUniqueId, created on rank 0 and handed to the other ranks by the caller."""

    _fields_ = [("bytes", C.c_ubyte * 128)]


def group_unique_id() -> bytes:
    gid = GroupId()
    rc = load().fn("group_unique_id")(C.byref(gid))
    if rc:
        raise Ag2Error(rc, "ag2_group_unique_id")
    return bytes(gid.bytes)


class Group:
    """ag2_group: the path over several GPUs behind the C ABI (csrc/group.cu).

    Group(params, n_gpus=N)                      one process, N host threads inside the library
    Group(params, uid=..., rank=r, world=w, device=d)   one process per rank (torchrun); collective
    """

    def __init__(self, params: Params, n_gpus: int = 0, uid: Optional[bytes] = None, rank: int = 0, world: int = 1,
                 device: int = 0):
        self.api = load()
        self._h = C.c_void_p()
        self.params = params.copy()
        if n_gpus:
            rc = self.api.fn("group_create")(C.byref(self.params), C.c_int(n_gpus), C.byref(self._h))
            self.rank, self.world = 0, n_gpus
        else:
            gid = GroupId()
            if uid is not None:
                C.memmove(C.byref(gid), uid, min(len(uid), 128))
            rc = self.api.fn("group_create_rank")(C.byref(self.params), C.byref(gid) if world > 1 else None,
                                                  C.c_int(rank), C.c_int(world), C.c_int(device), C.byref(self._h))
            self.rank, self.world = rank, world
        if rc:
            raise Ag2Error(rc, "ag2_group_create")
        w, nl = C.c_int(), C.c_int()
        self.api.fn("group_size")(self._h, C.byref(w), C.byref(nl))
        self.n_local = int(nl.value)
        self._out = None

    def close(self):
        if self._h:
            self.api.fn("group_destroy")(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc: int, what: str):
        if rc:
            f = self.api.lib.ag2_group_last_error
            f.restype = C.c_char_p
            raise Ag2Error(rc, what, (f(self._h) or b"").decode())

    def local_ctx(self, i: int = 0) -> Context:
        h = C.c_void_p()
        self._check(self.api.fn("group_local_ctx")(self._h, C.c_int(i), C.byref(h)), "ag2_group_local_ctx")
        return _BorrowedContext(h, self.params, self.api)

    def init_lenet_xavier(self, seed: int):
        for i in range(self.n_local):
            self.local_ctx(i).init_lenet_xavier(seed)

    def set_lenet_weights(self, weights: dict):
        for i in range(self.n_local):
            self.local_ctx(i).set_lenet_weights(weights)

    def peer_stores(self) -> bool:
        v = C.c_int()
        self._check(self.api.fn("group_transport")(self._h, C.byref(v)), "ag2_group_transport")
        return bool(v.value)

    def _result_buffer(self, cap: int) -> np.ndarray:
        if self._out is None or len(self._out) < cap:
            self._out = np.zeros(max(cap, 1), dtype=HYPOTHESIS_DTYPE)
        return self._out

    def _cap(self, n_samples: int, rounds: int = 0, per_round: int = 0) -> int:
        p = self.params
        if p.num_selected > 0 and p.antipodal_mode != MODE_NONE:
            return int(p.num_selected) * (rounds + 1)
        return (n_samples + rounds * per_round) * int(p.num_orientations)

    def detect_ptr(self, xyz_ptr: int, n: int, cam_ptr: int, num_cams: int, smp_ptr: int, n_samples: int,
                   on_device: bool) -> np.ndarray:
        """Raw-pointer form (host pointers, or device pointers on rank 0's GPU); pointers are read on rank 0 only."""
        out = self._result_buffer(self._cap(n_samples))
        got = C.c_size_t()
        rc = self.api.fn("group_detect")(self._h, C.c_void_p(xyz_ptr or None), C.c_size_t(n), C.c_void_p(cam_ptr or None),
                                         C.c_int(num_cams), C.c_void_p(smp_ptr or None), C.c_size_t(n_samples),
                                         C.c_int(int(on_device)), _fptr(out), C.c_size_t(len(out)), C.byref(got))
        self._check(rc, "ag2_group_detect")
        return out[: int(got.value)].copy()

    def io_bytes(self):
        """(host->device, device->host) bytes the first local rank moved in the last call."""
        a, b = C.c_size_t(), C.c_size_t()
        self._check(self.api.fn("group_io_bytes")(self._h, C.byref(a), C.byref(b)), "ag2_group_io_bytes")
        return int(a.value), int(b.value)

    def last_timing(self) -> dict:
        """Per-segment account of the first local rank's last detect call (ag2_group_last_timing), ms."""
        a = (C.c_double * 10)()
        self._check(self.api.fn("group_last_timing")(self._h, a), "ag2_group_last_timing")
        names = ["inputs", "grid", "normals_own", "normals_store", "normals_wait", "detect", "results_wait",
                 "host_exchange", "host_total", "device_total"]
        return {k: round(float(v), 4) for k, v in zip(names, a)}

    def detect_importance_ptr(self, xyz_ptr: int, n: int, cam_ptr: int, num_cams: int, smp_ptr: int, n_samples: int,
                              on_device: bool, ip: IsParams) -> np.ndarray:
        out = self._result_buffer(self._cap(int(n_samples), int(ip.num_iterations), int(ip.num_samples)))
        got = C.c_size_t()
        rc = self.api.fn("group_detect_importance")(
            self._h, C.c_void_p(xyz_ptr or None), C.c_size_t(int(n)), C.c_void_p(cam_ptr or None), C.c_int(num_cams),
            C.c_void_p(smp_ptr or None), C.c_size_t(int(n_samples)), C.c_int(int(on_device)), C.byref(ip), _fptr(out),
            C.c_size_t(len(out)), C.byref(got))
        self._check(rc, "ag2_group_detect_importance")
        return out[: int(got.value)].copy()

    def detect(self, xyz, cam_mask, num_cams: int, samples, n: Optional[int] = None, n_samples: Optional[int] = None):
        """numpy inputs on rank 0 (None elsewhere, with n / n_samples given)."""
        if xyz is not None:
            xyz = np.ascontiguousarray(xyz, dtype=np.float32).reshape(-1, 3)
            cam = None if cam_mask is None else np.ascontiguousarray(cam_mask, dtype=np.uint8)
            smp = np.ascontiguousarray(samples, dtype=np.int32)
            self._keep = (xyz, cam, smp)
            return self.detect_ptr(xyz.ctypes.data, len(xyz), 0 if cam is None else cam.ctypes.data, num_cams,
                                   smp.ctypes.data, len(smp), False)
        return self.detect_ptr(0, int(n), 0, num_cams, 0, int(n_samples), False)

    def detect_importance(self, xyz, cam_mask, num_cams: int, samples, ip: IsParams, n: Optional[int] = None,
                          n_samples: Optional[int] = None) -> np.ndarray:
        if xyz is not None:
            xyz = np.ascontiguousarray(xyz, dtype=np.float32).reshape(-1, 3)
            cam = None if cam_mask is None else np.ascontiguousarray(cam_mask, dtype=np.uint8)
            smp = np.ascontiguousarray(samples, dtype=np.int32)
            n, n_samples = len(xyz), len(smp)
            ptrs = (xyz.ctypes.data, 0 if cam is None else cam.ctypes.data, smp.ctypes.data)
        else:
            ptrs = (0, 0, 0)
        self._keep = (xyz, cam_mask, samples)
        return self.detect_importance_ptr(ptrs[0], int(n), ptrs[1], num_cams, ptrs[2], int(n_samples), False, ip)
