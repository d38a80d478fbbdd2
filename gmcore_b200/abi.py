"""ctypes mirror of include/gmcore_gpu.h (gmc_config, gmc_mesh_vectors).

Field order and types must match the header exactly; tests/test_abi.py checks
the struct sizes against the compiled library.
"""
import ctypes as C

GMC_EULER, GMC_PC2, GMC_WRFRK3 = 0, 1, 2
GMC_PV_MIDPOINT, GMC_PV_UPWIND, GMC_PV_TVD = 0, 1, 2
GMC_TVD_NONE, GMC_TVD_VAN_LEER, GMC_TVD_MC = 0, 1, 2
GMC_BIG_FILTER, GMC_SMALL_FILTER1, GMC_SMALL_FILTER2 = 0, 1, 2
GMC_FORCING_NONE, GMC_FORCING_HELD_SUAREZ = 0, 1
GMC_LIMITER_NONE, GMC_LIMITER_MONO, GMC_LIMITER_PD = 0, 1, 2
GMC_FLUX_PPM, GMC_FLUX_VAN_LEER = 0, 1

_INT_FIELDS = ["nlon", "nlat", "nlev", "hx", "hy", "lat_ibeg", "lat_iend", "device", "rank", "nranks",
               "baroclinic", "time_scheme", "ke_scheme", "pv_scheme", "upwind_order_pv", "pv_pole_stokes",
               "coriolis_scheme", "use_div_damp", "div_damp_order", "div_damp_k0", "use_smag_damp",
               "test_forcing", "ntracers", "limiter_type", "ffsl_flux_type", "tvd_limiter_type"]
_DBL_FIELDS = ["ke_cell_wgt", "upwind_wgt_pv", "div_damp_coef2", "div_damp_coef4", "div_damp_top",
               "div_damp_pole", "div_damp_lat0", "smag_damp_coef", "dt_dyn", "dt_adv",
               "omega", "radius", "g", "Rd", "cpd", "p0", "Rd_o_cpd", "ptop", "hybrid_p0"]


class GmcConfig(C.Structure):
    _fields_ = [(n, C.c_int) for n in _INT_FIELDS] + [(n, C.c_double) for n in _DBL_FIELDS]


MESH_VECTOR_NAMES = [
    "full_lat", "half_lat", "full_cos_lat", "half_cos_lat", "full_sin_lat", "half_sin_lat",
    "area_cell", "area_lon", "area_lon_west", "area_lon_east", "area_lon_north", "area_lon_south",
    "area_lat", "area_lat_west", "area_lat_east", "area_lat_north", "area_lat_south", "area_vtx",
    "area_subcell", "de_lon", "le_lon", "de_lat", "le_lat", "full_f", "half_f",
    "full_tangent_wgt", "half_tangent_wgt", "half_lon",
    "hyai", "hybi", "full_lev", "half_lev", "full_dlev", "half_dlev"]


class GmcMeshVectors(C.Structure):
    _fields_ = [(n, C.POINTER(C.c_double)) for n in MESH_VECTOR_NAMES]


def dptr(a):
    """double* of a C-contiguous float64 numpy array (None -> NULL)."""
    if a is None:
        return C.POINTER(C.c_double)()
    assert a.dtype.name == "float64" and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_double))


def iptr(a):
    assert a.dtype.name == "int32" and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_int))
