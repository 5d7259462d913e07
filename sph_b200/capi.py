"""ctypes binding of libsphb200.so (include/sphb200.h) -- the only way Python reaches the engine.

There is no Python or CPU implementation behind these calls: if the shared library is missing or a
call fails, an exception is raised.  The reference-side equivalent of this file is the cgo/ctypes
stub shown in INTEGRATION.md.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SPHB_LIB") or os.path.join(_HERE, "libsphb200.so")   # SPHB_LIB: an experimental build of the same library

# Host exchange records, byte-compatible with the reference:
# SPH/Particle.hpp:42-50 (52 B), SPH/Wall.hpp:28-31 (32 B), SPH/Sphere.hpp:29-33 (44 B)
PARTICLE = np.dtype(
    [("pos", "<f4", 3), ("vel", "<f4", 3), ("accel", "<f4", 3), ("mass", "<f4"), ("density", "<f4"),
     ("pressure", "<f4"), ("thermal", "<f4")]
)
WALL = np.dtype([("normal", "<f4", 3), ("center", "<f4", 3), ("xlength", "<f4"), ("ylength", "<f4")])
SPHERE = np.dtype([("pos", "<f4", 3), ("vel", "<f4", 3), ("accel", "<f4", 3), ("radius", "<f4"), ("mass", "<f4")])

F_POS, F_VEL, F_ACCEL, F_DENSITY, F_ALL = 1, 2, 4, 8, 15
DOMAIN_BBOX, DOMAIN_FIXED = 0, 1
NOALLOC_MAX_POINTS = 4194304   # SPHB_NOALLOC_MAX_POINTS
MG_EXPORT_BYTES = 64            # SPHB_MG_EXPORT_BYTES


class SphbError(RuntimeError):
    pass


class Camera(C.Structure):
    """struct sphb_camera (include/sphb200.h)."""

    _fields_ = [("mvp", C.c_float * 16), ("width", C.c_int32), ("height", C.c_int32), ("point_scale", C.c_float),
                ("sprite_radius", C.c_float), ("clear_value", C.c_float), ("depth_test", C.c_int32)]


class Slab(C.Structure):
    """struct sphb_slab (include/sphb200.h)."""

    _fields_ = [("lo", C.c_float), ("hi", C.c_float), ("halo", C.c_float), ("has_left", C.c_int32), ("has_right", C.c_int32),
                ("shift_left", C.c_float), ("shift_right", C.c_float), ("periodic_y", C.c_int32), ("periodic_z", C.c_int32),
                ("ylo", C.c_float), ("ylen", C.c_float), ("zlo", C.c_float), ("zlen", C.c_float)]


class Stats(C.Structure):
    """struct sphb_stats (include/sphb200.h)."""

    _fields_ = [("arrival_order_cells", C.c_uint64), ("steps", C.c_uint64), ("launches", C.c_uint64), ("particles", C.c_uint64),
                ("capacity", C.c_uint64), ("reserved", C.c_uint64 * 3)]


class Params(C.Structure):
    """struct sphb_params (include/sphb200.h)."""

    _fields_ = [
        ("h", C.c_float), ("rho0", C.c_float), ("c", C.c_float), ("dt", C.c_float), ("eps", C.c_float),
        ("eos_k", C.c_float), ("gravity", C.c_float * 3),
        ("visc_alpha", C.c_float), ("visc_beta", C.c_float), ("visc_eta", C.c_float),
        ("restitution", C.c_float), ("cell_size", C.c_float),
        ("wall_test_off", C.c_double), ("wall_test_lim", C.c_double),
        ("domain_mode", C.c_int32), ("domain_lo", C.c_float * 3), ("domain_hi", C.c_float * 3),
        ("periodic_reserved", C.c_int32 * 3), ("max_cells", C.c_int64), ("write_accel", C.c_int32),
        ("cell_x_div", C.c_int32), ("laplacian_viscosity", C.c_int32), ("lap_visc_coef", C.c_float), ("parity_mode", C.c_int32), ("reserved", C.c_int32 * 3),
    ]


EXPORTS = [
    "sphb_default_params", "sphb_last_error", "sphb_version", "sphb_params_size", "sphb_create", "sphb_destroy", "sphb_set_walls",
    "sphb_upload_aos", "sphb_download_aos", "sphb_update_cells", "sphb_update_density", "sphb_update_forces",
    "sphb_update_posvel", "sphb_collision", "sphb_step", "sphb_step_aos", "sphb_synchronize",
    "sphb_host_register", "sphb_host_unregister", "sphb_eval_kernels", "sphb_get_cells", "sphb_get_bbox",
    "sphb_count", "sphb_enable_timing", "sphb_get_timing", "sphb_launch_count", "sphb_upload_soa_device",
    "sphb_get_device_state", "sphb_stream", "sphb_bbox_noalloc", "sphb_download_soa",
    "sphb_count_nonfinite", "sphb_default_camera", "sphb_render_depth", "sphb_download_positions", "sphb_set_spheres", "sphb_get_spheres", "sphb_mg_message_bytes", "sphb_reserve", "sphb_mg_pack", "sphb_mg_pack_limits", "sphb_mg_unpack", "sphb_mg_sent", "sphb_get_stats", "sphb_mg_unpack_async", "sphb_mg_poll", "sphb_mg_link_create", "sphb_mg_link_destroy", "sphb_mg_link_export",
    "sphb_mg_link_connect_ipc", "sphb_mg_link_connect_local", "sphb_mg_exchange", "sphb_set_ids", "sphb_multi_create", "sphb_multi_destroy", "sphb_multi_last_error", "sphb_multi_set_walls",
    "sphb_multi_upload_aos", "sphb_multi_step", "sphb_multi_synchronize", "sphb_multi_download_aos", "sphb_multi_step_aos", "sphb_multi_info", "sphb_multi_plan",
]

_lib = None


def load_library() -> C.CDLL:
    """Load libsphb200.so.  Fails loudly: there is no fallback implementation."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SphbError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C sph_b200/csrc`).  sph_b200 has no CPU or PyTorch fallback."
        )
    L = C.CDLL(LIB_PATH)
    vp, sz, u32, i32 = C.c_void_p, C.c_size_t, C.c_uint32, C.c_int
    L.sphb_default_params.argtypes = [C.POINTER(Params)]
    L.sphb_default_params.restype = None
    L.sphb_last_error.restype = C.c_char_p
    L.sphb_version.restype = C.c_char_p
    L.sphb_params_size.restype = sz
    if L.sphb_params_size() != C.sizeof(Params):
        raise SphbError(f"sphb_params layout mismatch: library {L.sphb_params_size()} B, binding {C.sizeof(Params)} B")
    L.sphb_create.argtypes = [C.POINTER(Params), i32, C.POINTER(vp)]
    L.sphb_destroy.argtypes = [vp]
    L.sphb_destroy.restype = None
    L.sphb_set_walls.argtypes = [vp, vp, sz, sz]
    L.sphb_set_spheres.argtypes = [vp, vp, sz, sz]
    L.sphb_get_spheres.argtypes = [vp, vp, sz, sz]
    L.sphb_upload_aos.argtypes = [vp, vp, sz, sz]
    L.sphb_download_aos.argtypes = [vp, vp, sz, sz, u32]
    L.sphb_download_positions.argtypes = [vp, vp, sz]
    L.sphb_default_camera.argtypes = [C.POINTER(Camera)]
    L.sphb_default_camera.restype = None
    L.sphb_render_depth.argtypes = [vp, C.POINTER(Camera), vp]
    for name in ("sphb_update_cells", "sphb_update_density", "sphb_update_forces", "sphb_update_posvel",
                 "sphb_collision", "sphb_synchronize"):
        getattr(L, name).argtypes = [vp]
    L.sphb_step.argtypes = [vp, i32]
    L.sphb_step_aos.argtypes = [vp, vp, sz, sz, i32, u32]
    L.sphb_host_register.argtypes = [vp, sz]
    L.sphb_host_unregister.argtypes = [vp]
    L.sphb_eval_kernels.argtypes = [vp, vp, vp, sz, vp, vp, vp]
    L.sphb_get_cells.argtypes = [vp, vp, vp, vp, vp, vp]
    L.sphb_get_bbox.argtypes = [vp, vp]
    L.sphb_count.argtypes = [vp]
    L.sphb_count.restype = sz
    L.sphb_count_nonfinite.argtypes = [vp, C.POINTER(sz)]
    L.sphb_enable_timing.argtypes = [vp, i32]
    L.sphb_get_timing.argtypes = [vp, vp, vp]
    L.sphb_launch_count.argtypes = [vp]
    L.sphb_launch_count.restype = C.c_uint64
    L.sphb_upload_soa_device.argtypes = [vp, vp, vp, vp, sz]
    L.sphb_get_device_state.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    L.sphb_stream.argtypes = [vp]
    L.sphb_stream.restype = vp
    L.sphb_download_soa.argtypes = [vp, vp, vp, vp, vp, vp]
    L.sphb_bbox_noalloc.argtypes = [vp, sz, vp, vp]
    L.sphb_mg_message_bytes.argtypes = [sz]
    L.sphb_mg_message_bytes.restype = sz
    L.sphb_reserve.argtypes = [vp, sz]
    L.sphb_mg_pack.argtypes = [vp, C.POINTER(Slab), vp, vp, sz]
    L.sphb_mg_unpack.argtypes = [vp, C.POINTER(Slab), vp, vp, C.POINTER(sz * 4)]
    L.sphb_mg_sent.argtypes = [vp, C.POINTER(sz * 2)]
    L.sphb_mg_pack_limits.argtypes = [vp, C.POINTER(Slab), vp, sz, vp, sz]
    L.sphb_get_stats.argtypes = [vp, C.POINTER(Stats)]
    L.sphb_mg_unpack_async.argtypes = [vp, C.POINTER(Slab), vp, vp]
    L.sphb_mg_poll.argtypes = [vp, C.POINTER(sz * 8), i32]
    L.sphb_mg_link_create.argtypes = [vp, sz, C.POINTER(vp)]
    L.sphb_mg_link_destroy.argtypes = [vp]
    L.sphb_mg_link_destroy.restype = None
    L.sphb_mg_link_export.argtypes = [vp, C.c_char_p]
    L.sphb_mg_link_connect_ipc.argtypes = [vp, C.c_char_p, C.c_char_p]
    L.sphb_mg_link_connect_local.argtypes = [vp, vp, vp]
    L.sphb_mg_exchange.argtypes = [vp, vp, C.POINTER(Slab)]
    L.sphb_set_ids.argtypes = [vp, vp, sz]
    L.sphb_multi_create.argtypes = [i32, vp, C.POINTER(Params), C.POINTER(vp)]
    L.sphb_multi_destroy.argtypes = [vp]
    L.sphb_multi_destroy.restype = None
    L.sphb_multi_last_error.restype = C.c_char_p
    L.sphb_multi_set_walls.argtypes = [vp, vp, sz, sz]
    L.sphb_multi_upload_aos.argtypes = [vp, vp, sz, sz]
    L.sphb_multi_step.argtypes = [vp, i32]
    L.sphb_multi_synchronize.argtypes = [vp]
    L.sphb_multi_download_aos.argtypes = [vp, vp, sz, sz, u32]
    L.sphb_multi_step_aos.argtypes = [vp, vp, sz, sz, i32, u32]
    L.sphb_multi_info.argtypes = [vp, C.POINTER(i32), vp, vp]
    L.sphb_multi_plan.argtypes = [vp, sz, i32, C.c_float, vp]
    _lib = L
    return L


def default_params(**overrides) -> Params:
    L = load_library()
    p = Params()
    L.sphb_default_params(C.byref(p))
    for k, v in overrides.items():
        cur = getattr(p, k)
        if isinstance(cur, C.Array):
            for i, x in enumerate(v):
                cur[i] = x
        else:
            setattr(p, k, v)
    return p


def default_camera() -> Camera:
    cam = Camera()
    load_library().sphb_default_camera(C.byref(cam))
    return cam


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    return a.ctypes.data_as(C.c_void_p)


def multi_plan(x: np.ndarray, ndev: int, h: float = 0.45):
    """The slab cuts sphb_multi_upload_aos would choose for these x coordinates (host arithmetic, no GPU)."""
    L = load_library()
    x = np.ascontiguousarray(x, np.float32)
    b = np.zeros(ndev + 1, np.float64)
    rc = L.sphb_multi_plan(_ptr(x), len(x), ndev, C.c_float(h), _ptr(b))
    if rc:
        raise SphbError(f"sphb error {rc}: {L.sphb_multi_last_error().decode()}")
    return b


class Multi:
    """sphb_multi_*: several GPUs of one box behind one handle, driven by this one host thread."""

    def __init__(self, devices, params: "Params | None" = None, **overrides):
        self.lib = load_library()
        self.params = params if params is not None else default_params(**overrides)
        devs = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        self._check(self.lib.sphb_multi_create(len(devices), devs, C.byref(self.params), C.byref(h)))
        self.h = h
        self.ndev = len(devices)

    def _check(self, rc: int):
        if rc != 0:
            raise SphbError(f"sphb error {rc}: {self.lib.sphb_multi_last_error().decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.lib.sphb_multi_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_walls(self, walls: np.ndarray):
        walls = np.ascontiguousarray(walls)
        assert walls.dtype == WALL
        self._check(self.lib.sphb_multi_set_walls(self.h, _ptr(walls), len(walls), WALL.itemsize))

    def upload(self, particles: np.ndarray):
        assert particles.dtype == PARTICLE and particles.flags.c_contiguous
        self._check(self.lib.sphb_multi_upload_aos(self.h, _ptr(particles), len(particles), PARTICLE.itemsize))

    def step(self, nsteps: int = 1):
        self._check(self.lib.sphb_multi_step(self.h, nsteps))

    def synchronize(self):
        self._check(self.lib.sphb_multi_synchronize(self.h))

    def download(self, particles: np.ndarray, mask: int = F_ALL):
        assert particles.dtype == PARTICLE and particles.flags.c_contiguous
        self._check(self.lib.sphb_multi_download_aos(self.h, _ptr(particles), len(particles), PARTICLE.itemsize, mask))
        return particles

    def info(self):
        counts = (C.c_size_t * self.ndev)()
        bounds = (C.c_double * (self.ndev + 1))()
        nd = C.c_int(0)
        self._check(self.lib.sphb_multi_info(self.h, C.byref(nd), counts, bounds))
        return [int(c) for c in counts], [float(b) for b in bounds]


class Engine:
    """Thin object wrapper over the C ABI; one instance = one sphb_engine on one GPU."""

    def __init__(self, params: Params | None = None, device: int = 0, **overrides):
        self.lib = load_library()
        self.params = params if params is not None else default_params(**overrides)
        h = C.c_void_p()
        self._check(self.lib.sphb_create(C.byref(self.params), device, C.byref(h)))
        self.h = h
        self.device = device

    def _check(self, rc: int):
        if rc != 0:
            raise SphbError(f"sphb error {rc}: {self.lib.sphb_last_error().decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.lib.sphb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # --- state
    def set_walls(self, walls: np.ndarray):
        walls = np.ascontiguousarray(walls)
        if walls.dtype == WALL:
            self._check(self.lib.sphb_set_walls(self.h, _ptr(walls), len(walls), WALL.itemsize))
        else:
            w = np.ascontiguousarray(walls, np.float32).reshape(-1, 6)
            self._check(self.lib.sphb_set_walls(self.h, _ptr(w), len(w), 24))

    def set_spheres(self, spheres: np.ndarray | None):
        if spheres is None or len(spheres) == 0:
            self._check(self.lib.sphb_set_spheres(self.h, None, 0, SPHERE.itemsize))
            return
        assert spheres.dtype == SPHERE and spheres.flags.c_contiguous
        self._check(self.lib.sphb_set_spheres(self.h, _ptr(spheres), len(spheres), SPHERE.itemsize))

    def get_spheres(self, spheres: np.ndarray):
        assert spheres.dtype == SPHERE and spheres.flags.c_contiguous
        self._check(self.lib.sphb_get_spheres(self.h, _ptr(spheres), len(spheres), SPHERE.itemsize))
        return spheres

    def upload(self, particles: np.ndarray):
        assert particles.dtype == PARTICLE and particles.flags.c_contiguous
        self._check(self.lib.sphb_upload_aos(self.h, _ptr(particles), len(particles), PARTICLE.itemsize))

    def download(self, particles: np.ndarray, mask: int = F_ALL):
        assert particles.dtype == PARTICLE and particles.flags.c_contiguous
        self._check(self.lib.sphb_download_aos(self.h, _ptr(particles), len(particles), PARTICLE.itemsize, mask))
        return particles

    def render_depth(self, camera: "Camera | None" = None) -> np.ndarray:
        """The viewer's picture of the current state as a float image [height, width], row 0 at the bottom."""
        cam = camera if camera is not None else default_camera()
        img = np.zeros((cam.height, cam.width), np.float32)
        self._check(self.lib.sphb_render_depth(self.h, C.byref(cam), _ptr(img)))
        return img

    def download_positions(self, xyz: np.ndarray | None = None, ptr: int | None = None, n: int | None = None):
        """Packed x,y,z in original order into `xyz` (float32 [n,3]) or a raw host pointer."""
        if ptr is not None:
            self._check(self.lib.sphb_download_positions(self.h, C.c_void_p(ptr), n))
            return None
        if xyz is None:
            xyz = np.zeros((self.count(), 3), np.float32)
        self._check(self.lib.sphb_download_positions(self.h, _ptr(xyz), len(xyz)))
        return xyz

    def download_first(self, particles: np.ndarray, mask: int = F_ALL):
        """Download the first len(particles) particles of the original order (the rest may be ghosts)."""
        return self.download(particles, mask)

    def download_soa(self):
        """(posm, vel, aux, accel, ids) of the current state in SORTED order."""
        n = self.count()
        arrs = [np.zeros((n, 4), np.float32) for _ in range(4)] + [np.zeros(n, np.int32)]
        self._check(self.lib.sphb_download_soa(self.h, *[_ptr(a) for a in arrs]))
        return tuple(arrs)

    def upload_ptr(self, ptr: int, n: int, stride: int = 52):
        self._check(self.lib.sphb_upload_aos(self.h, C.c_void_p(ptr), n, stride))

    def download_ptr(self, ptr: int, n: int, stride: int = 52, mask: int = F_ALL):
        self._check(self.lib.sphb_download_aos(self.h, C.c_void_p(ptr), n, stride, mask))

    def upload_soa_device(self, posm_ptr: int, vel_ptr: int, n: int, ids_ptr: int | None = None):
        self._check(self.lib.sphb_upload_soa_device(self.h, C.c_void_p(posm_ptr), C.c_void_p(vel_ptr),
                                                    C.c_void_p(ids_ptr) if ids_ptr else None, n))

    def set_ids(self, ids: np.ndarray):
        """Caller's labels (int32, upload order); they travel with the particles through slab exchanges."""
        ids = np.ascontiguousarray(ids, np.int32)
        self._check(self.lib.sphb_set_ids(self.h, _ptr(ids), len(ids)))

    # --- multi-GPU slab decomposition
    def reserve(self, n_max: int):
        self._check(self.lib.sphb_reserve(self.h, n_max))

    def mg_pack(self, slab: Slab, msg_left: int | None, msg_right: int | None, capacity: int):
        self._check(self.lib.sphb_mg_pack(self.h, C.byref(slab), C.c_void_p(msg_left) if msg_left else None,
                                          C.c_void_p(msg_right) if msg_right else None, capacity))

    def mg_pack_limits(self, slab: Slab, msg_left: int | None, cap_left: int, msg_right: int | None, cap_right: int):
        self._check(self.lib.sphb_mg_pack_limits(self.h, C.byref(slab), C.c_void_p(msg_left) if msg_left else None, cap_left,
                                                 C.c_void_p(msg_right) if msg_right else None, cap_right))

    def mg_unpack(self, slab: Slab, from_left: int | None, from_right: int | None):
        counts = (C.c_size_t * 4)()
        self._check(self.lib.sphb_mg_unpack(self.h, C.byref(slab), C.c_void_p(from_left) if from_left else None,
                                            C.c_void_p(from_right) if from_right else None, C.byref(counts)))
        return tuple(int(c) for c in counts)

    def mg_unpack_async(self, slab: Slab, from_left: int | None, from_right: int | None):
        self._check(self.lib.sphb_mg_unpack_async(self.h, C.byref(slab), C.c_void_p(from_left) if from_left else None,
                                                  C.c_void_p(from_right) if from_right else None))

    def mg_poll(self, wait: bool = True):
        """(kept, from_left, from_right, total, dead, sent_left, sent_right, reports so far) of the latest asynchronous unpack."""
        st = (C.c_size_t * 8)()
        self._check(self.lib.sphb_mg_poll(self.h, C.byref(st), int(wait)))
        return tuple(int(x) for x in st)

    # --- peer-to-peer halo link
    def mg_link_create(self, capacity: int) -> int:
        h = C.c_void_p()
        self._check(self.lib.sphb_mg_link_create(self.h, capacity, C.byref(h)))
        return h.value

    def mg_link_destroy(self, link: int):
        self.lib.sphb_mg_link_destroy(C.c_void_p(link))

    def mg_link_export(self, link: int) -> bytes:
        buf = C.create_string_buffer(MG_EXPORT_BYTES)
        self._check(self.lib.sphb_mg_link_export(C.c_void_p(link), buf))
        return buf.raw

    def mg_link_connect_ipc(self, link: int, left: bytes | None, right: bytes | None):
        self._check(self.lib.sphb_mg_link_connect_ipc(C.c_void_p(link), left, right))

    def mg_link_connect_local(self, link: int, left: int | None, right: int | None):
        self._check(self.lib.sphb_mg_link_connect_local(C.c_void_p(link), C.c_void_p(left) if left else None, C.c_void_p(right) if right else None))

    def mg_exchange(self, link: int, slab: Slab):
        self._check(self.lib.sphb_mg_exchange(self.h, C.c_void_p(link), C.byref(slab)))

    def mg_sent(self):
        sent = (C.c_size_t * 2)()
        self._check(self.lib.sphb_mg_sent(self.h, C.byref(sent)))
        return int(sent[0]), int(sent[1])

    # --- the SphUtils methods
    def update_cells(self):
        self._check(self.lib.sphb_update_cells(self.h))

    def update_density(self):
        self._check(self.lib.sphb_update_density(self.h))

    def update_forces(self):
        self._check(self.lib.sphb_update_forces(self.h))

    def update_posvel(self):
        self._check(self.lib.sphb_update_posvel(self.h))

    def collision(self):
        self._check(self.lib.sphb_collision(self.h))

    def step(self, nsteps: int = 1):
        self._check(self.lib.sphb_step(self.h, nsteps))

    def step_aos(self, particles: np.ndarray, nsteps: int = 1, mask: int = F_ALL):
        assert particles.dtype == PARTICLE and particles.flags.c_contiguous
        self._check(self.lib.sphb_step_aos(self.h, _ptr(particles), len(particles), PARTICLE.itemsize, nsteps, mask))

    def step_aos_ptr(self, ptr: int, n: int, nsteps: int = 1, mask: int = F_ALL, stride: int = 52):
        self._check(self.lib.sphb_step_aos(self.h, C.c_void_p(ptr), n, stride, nsteps, mask))

    def synchronize(self):
        self._check(self.lib.sphb_synchronize(self.h))

    # --- diagnostics
    def eval_kernels(self, xi, xj):
        xi = np.ascontiguousarray(xi, np.float32).reshape(-1, 3)
        xj = np.ascontiguousarray(xj, np.float32).reshape(-1, 3)
        n = len(xi)
        w = np.zeros(n, np.float32)
        g = np.zeros((n, 3), np.float32)
        lap = np.zeros(n, np.float32)
        self._check(self.lib.sphb_eval_kernels(self.h, _ptr(xi), _ptr(xj), n, _ptr(w), _ptr(g), _ptr(lap)))
        return w, g, lap

    def get_cells(self):
        n = self.count()
        xyz = np.zeros((n, 3), np.int32)
        perm = np.zeros(n, np.int32)
        dims = np.zeros(3, np.int32)
        origin = np.zeros(3, np.float32)
        cell = np.zeros(3, np.float32)
        self._check(self.lib.sphb_get_cells(self.h, _ptr(xyz), _ptr(perm), _ptr(dims), _ptr(origin), _ptr(cell)))
        return xyz, perm, dims, origin, cell

    def get_bbox(self) -> np.ndarray:
        b = np.zeros(6, np.float32)
        self._check(self.lib.sphb_get_bbox(self.h, _ptr(b)))
        return b

    def count(self) -> int:
        return int(self.lib.sphb_count(self.h))

    def count_nonfinite(self) -> int:
        k = C.c_size_t(0)
        self._check(self.lib.sphb_count_nonfinite(self.h, C.byref(k)))
        return int(k.value)

    def enable_timing(self, on: bool = True):
        self._check(self.lib.sphb_enable_timing(self.h, int(on)))

    def get_timing(self):
        """(ms[4] averaged per step: sort, density, force+integrate+collide, whole step; steps averaged)."""
        t = np.zeros(4, np.float32)
        k = C.c_int(0)
        self._check(self.lib.sphb_get_timing(self.h, _ptr(t), C.byref(k)))
        return t, k.value

    def stats(self) -> dict:
        st = Stats()
        self._check(self.lib.sphb_get_stats(self.h, C.byref(st)))
        return {k: int(getattr(st, k)) for k, _ in Stats._fields_ if k != "reserved"}

    def launch_count(self) -> int:
        return int(self.lib.sphb_launch_count(self.h))

    def device_state(self):
        ps = [C.c_void_p() for _ in range(5)]
        self._check(self.lib.sphb_get_device_state(self.h, *[C.byref(p) for p in ps]))
        return tuple(p.value for p in ps)

    def stream(self) -> int:
        return int(self.lib.sphb_stream(self.h) or 0)
