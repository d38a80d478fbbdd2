/* gmcore_host.h — C ABI of the stand-alone host set-up (mesh geometry, filter
 * weights, vertical coordinate, analytic initial conditions).  This is NOT the hot
 * path: in production GMCORE's Fortran drivers own all of this and pass the
 * results through include/gmcore_gpu.h.  It exists because no Fortran toolchain
 * is available on the build/GPU machines, so bench.py and the tests need a host.
 * References: src/dynamics/mesh_mod.F90:134-447, filter_types_mod.F90:31-147,
 * vert_coord/hybrid_coord_mod.F90:34-171, the *_test_mod.F90 files under src/tests.
 */
#ifndef GMCORE_HOST_H
#define GMCORE_HOST_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct gmh_host gmh_host;

enum { GMH_IC_ROSSBY_HAURWITZ = 0, GMH_IC_MOUNTAIN_ZONAL_FLOW = 1, GMH_IC_BAROCLINIC_WAVE = 2,
       GMH_IC_HELD_SUAREZ = 3, GMH_IC_STEADY_GEOSTROPHIC = 4 };

/* process_create_blocks (process_mod.F90:196-268): builds the mesh with the
 * namelist halo (2), the three filters for dt_dyn, derives lon_halo_width from
 * big_filter%ngrid_lat, and re-initialises the mesh with it.  nlev==1 -> SWM;
 * nlev==32 uses template 'wrf_l32', otherwise 'equal_eta'. */
gmh_host* gmh_create(int nlon, int nlat, int nlev, double dt_dyn, double ptop);
/* + the mesh switches of namelist_mod.F90:85-86 (coarse_pole_mul /= 0: latitude spacing stretched towards the
 * poles, mesh_mod.F90:195-202) and tangent_wgt_scheme = 'thuburn09' (mesh_mod.F90:428-437). */
gmh_host* gmh_create_opts(int nlon, int nlat, int nlev, double dt_dyn, double ptop,
                          double coarse_pole_mul, double coarse_pole_decay, int tangent_wgt_thuburn09);
void gmh_destroy(gmh_host* h);
int gmh_hx(const gmh_host* h);

/* Named 1-D vectors: every name of gmc_mesh_vectors plus full_lat_deg, half_lat_deg,
 * full_lon, full_sin_lat ...  Returns the length (or -1); copies into buf if non-NULL. */
int gmh_get_vector(const gmh_host* h, const char* name, double* buf);

/* Filters: which = GMC_BIG_FILTER.. ; returns leading dimensions through ld[2]. */
int gmh_filter_dims(const gmh_host* h, int which, int* ld_lon, int* ld_lat);
int gmh_get_filter(const gmh_host* h, int which, int* ngrid_lon, double* wgt_lon, int* ngrid_lat, double* wgt_lat);

/* Initial conditions into halo'd Fortran-layout arrays (interior only is written).
 * gz/pt share one argument; phs and gzs are 2-D.  seed is used by Held–Suarez. */
int gmh_set_ic(const gmh_host* h, int test_case, double* u_lon, double* v_lat, double* pt_or_gz,
               double* phs, double* gzs, unsigned long long seed);

#ifdef __cplusplus
}
#endif
#endif
