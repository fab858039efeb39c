"""ctypes binding of libdg1d.so (include/dg1d.h).  No torch types cross this boundary.

The library is built in-tree by `make -C fem_b200/csrc` (or __graft_entry__.build()).  There is no
CPU fallback: if the shared object is missing, importing the compute layer raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get('DG1D_LIB') or os.path.join(_HERE, 'libdg1d.so')   # DG1D_LIB: A/B builds in one GPU session

PDE = {'linear': 0, 'burger': 1, 'euler': 2, 'varadv': 3}
FLUX = {'central': 0, 'upwind': 1, 'roe': 2, 'godunov': 3, 'rusanov': 4,
        # Euler only: dg1d/c/euler/src/dg.h:7-11 (flux.c), SURVEY 8f N4
        'lf': 5, 'ecusp': 6, 'hllc': 7, 'ausmdv': 8, 'lfc': 9}
EULER_FLUXES = ('rusanov', 'roe', 'lf', 'ecusp', 'hllc', 'ausmdv', 'lfc')
BC = {'periodic': 0, 'copy': 1, 'halo': 2,
      # scalar PDEs: the C++ twin's BCType (dg1d/scalar_legendre/dg.h:44); outflow is the copy face
      'inflow': 3, 'outflow': 1}
LIMITER = {'none': 0, 'tvb': 1}
GHOST = 4


class Dg1dError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [('pde', C.c_int32), ('flux', C.c_int32), ('degree', C.c_int32), ('limiter', C.c_int32),
                ('char_lim', C.c_int32), ('bc_left', C.c_int32), ('bc_right', C.c_int32), ('nprob', C.c_int32),
                ('ncell', C.c_int64), ('cell_offset', C.c_int64), ('ncell_global', C.c_int64),
                ('gamma', C.c_double)]


_dp = C.POINTER(C.c_double)
_vp = C.c_void_p

# name -> (restype, argtypes); exactly the declarations of include/dg1d.h
PROTOTYPES = {
    'dg1d_version': (C.c_int, []),
    'dg1d_last_error': (C.c_char_p, []),
    'dg1d_device_count': (C.c_int, []),
    'dg1d_create': (C.c_int, [C.POINTER(Config), C.POINTER(_vp)]),
    'dg1d_destroy': (None, [_vp]),
    'dg1d_set_operators': (C.c_int, [_vp, _dp, _dp, _dp, _dp, _dp, _dp, _dp, C.c_double, C.c_double]),
    'dg1d_set_mesh': (C.c_int, [_vp, C.c_double, C.c_double, _dp, C.c_double]),
    'dg1d_set_boundary': (C.c_int, [_vp, _vp, C.c_int64]),
    'dg1d_pitch': (C.c_int64, [_vp]),
    'dg1d_state_doubles': (C.c_int64, [_vp]),
    'dg1d_nvar': (C.c_int, [_vp]),
    'dg1d_pack_state': (C.c_int, [_vp, _vp, _vp, _vp]),
    'dg1d_unpack_state': (C.c_int, [_vp, _vp, _vp, _vp]),
    'dg1d_pack_cells': (C.c_int, [_vp, _vp, _vp, C.c_int64, C.c_int64, _vp]),
    'dg1d_unpack_cells': (C.c_int, [_vp, _vp, _vp, C.c_int64, C.c_int64, _vp]),
    'dg1d_project': (C.c_int, [_vp, _vp, _vp, _vp]),
    'dg1d_max_speed': (C.c_int, [_vp, _vp, _vp, _vp]),
    'dg1d_step_control': (C.c_int, [_vp, C.c_double, C.c_double, C.c_double, C.c_int64, _vp, _vp, _vp, _vp, _vp, _vp,
                                    _vp, C.c_int64, _vp]),
    'dg1d_stage': (C.c_int, [_vp, C.c_int, _vp, _vp, _vp, _vp, _vp, C.c_double, _vp, _vp]),
    'dg1d_stage_part': (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, C.c_double, _vp, _vp]),
    'dg1d_eval_points': (C.c_int, [_vp, _vp, _dp, C.c_int, _vp, _vp, _vp]),
    'dg1d_residual': (C.c_int, [_vp, _vp, _vp, _vp]),
    'dg1d_rk_update': (C.c_int, [_vp, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp]),
    'dg1d_limiter': (C.c_int, [_vp, _vp, _vp, C.c_double, _vp]),
    'dg1d_run_steps': (C.c_int, [_vp, C.c_int64, C.POINTER(_vp), C.c_double, C.c_double, C.c_double, C.c_int64,
                                 C.c_double, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int64, _vp]),
    'dg1d_halo_doubles': (C.c_int64, [_vp]),
    'dg1d_halo_pack': (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    'dg1d_halo_unpack': (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    'dg1d_dist_available': (C.c_int, []),
    'dg1d_dist_unique_id': (C.c_int, [C.c_char_p]),
    'dg1d_dist_create': (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(_vp)]),
    'dg1d_dist_destroy': (None, [_vp]),
    'dg1d_dist_invalidate_halo': (None, [_vp]),
    'dg1d_dist_reduce_speed': (C.c_int, [_vp, _vp, C.c_int, _vp]),
    'dg1d_dist_p2p_open': (C.c_int, [_vp, _vp, C.c_char_p]),
    'dg1d_dist_p2p_connect': (C.c_int, [_vp, C.c_char_p]),
    'dg1d_dist_use_nccl': (C.c_int, [_vp]),
    'dg1d_dist_transport': (C.c_int, [_vp]),
    'dg1d_dist_p2p_error': (C.c_int, [_vp, C.POINTER(C.c_int64)]),
    'dg1d_dist_run_steps': (C.c_int, [_vp, _vp, C.c_int64, C.POINTER(_vp), C.c_double, C.c_double, C.c_double, C.c_int64,
                                      C.c_double, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int64, _vp, _vp, _vp, _vp, _vp]),
    'dg1d_solve_host': (C.c_int, [_vp, _dp, C.c_double, C.c_double, C.c_double, C.c_int64, C.c_double, _dp,
                                  C.POINTER(C.c_int64)]),
    'dg1d_error_norm_scratch_doubles': (C.c_int64, [_vp]),
    'dg1d_error_norm': (C.c_int, [_vp, _vp, _vp, _dp, _dp, C.c_int, C.c_double, _vp, _vp, _vp]),
    'dg1d_error_norm_noalloc_max_doubles': (C.c_int64, []),
    'dg1d_error_norm_noalloc': (C.c_int, [_vp, _vp, _vp, _dp, _dp, C.c_int, C.c_double, _vp, _vp]),
    'dg1d_selftest_division': (C.c_int, [C.c_int64, C.c_uint64, C.c_int, C.POINTER(C.c_uint64)]),
    'dg1d_launch_count': (C.c_int64, []),
}

_lib = None


def load():
    """Loads libdg1d.so and types every exported symbol.  Raises Dg1dError if it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise Dg1dError('%s not found: build it with `make -C fem_b200/csrc` or __graft_entry__.build(); '
                            'there is no CPU fallback' % SO_PATH)
        L = C.CDLL(SO_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error():
    return load().dg1d_last_error().decode()


def check(rc):
    if rc != 0:
        msg = load().dg1d_last_error().decode()
        if rc == -1:
            raise ValueError(msg)
        raise Dg1dError('dg1d error %d: %s' % (rc, msg))
