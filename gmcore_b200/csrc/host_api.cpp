// C ABI of the stand-alone host set-up (include/gmcore_host.h).  Not the hot path.
#include "../../include/gmcore_host.h"
#include "../../include/gmcore_gpu.h"
#include "host_filter.hpp"
#include "host_vert.hpp"
#include <cstring>
#include <map>

using namespace gmh;

struct gmh_host {
  Consts c;
  Mesh mesh;
  VertCoord vc;
  FilterCfg fcfg;
  Filter filt[3];
  double ptop = 0;
  std::vector<double> hyai_v, hybi_v;
};

extern "C" {

gmh_host* gmh_create(int nlon, int nlat, int nlev, double dt_dyn, double ptop) {
  return gmh_create_opts(nlon, nlat, nlev, dt_dyn, ptop, 0.0, 100.0, 0);
}
gmh_host* gmh_create_opts(int nlon, int nlat, int nlev, double dt_dyn, double ptop,
                          double coarse_pole_mul, double coarse_pole_decay, int tangent_wgt_thuburn09) {
  gmh_host* h = new gmh_host();
  h->mesh.coarse_pole_mul = coarse_pole_mul; h->mesh.coarse_pole_decay = coarse_pole_decay;
  h->mesh.thuburn09 = tangent_wgt_thuburn09;
  h->ptop = ptop;
  h->fcfg.dt_dyn = dt_dyn;
  // gmcore_init: mesh with the namelist halo (2); block stage 1 builds the filters;
  // process_create_blocks derives lon_halo_width from big_filter%ngrid_lat; reinit.
  h->mesh.init(h->c, nlon, nlat, nlev, 2, 2);
  for (int t = 0; t < 3; ++t) filter_init(h->c, h->mesh, h->fcfg, t, h->filt[t]);
  int max_hw = 2;
  for (int j = 1; j <= h->mesh.num_half_lat; ++j) {
    int hw = (h->filt[0].nlat_at(j) - 1) / 2;
    if (hw > max_hw) max_hw = hw;
  }
  h->mesh.init(h->c, nlon, nlat, nlev, max_hw, 2);
  vert_coord_init(h->mesh, h->vc);
  return h;
}
void gmh_destroy(gmh_host* h) { delete h; }
int gmh_hx(const gmh_host* h) { return h->mesh.hx; }

int gmh_get_vector(const gmh_host* hc, const char* name, double* buf) {
  gmh_host* h = const_cast<gmh_host*>(hc);
  Mesh& m = h->mesh;
  std::map<std::string, const std::vector<double>*> v = {
    {"full_lat", &m.full_lat.d}, {"half_lat", &m.half_lat.d}, {"full_cos_lat", &m.full_cos_lat.d},
    {"half_cos_lat", &m.half_cos_lat.d}, {"full_sin_lat", &m.full_sin_lat.d}, {"half_sin_lat", &m.half_sin_lat.d},
    {"full_lat_deg", &m.full_lat_deg.d}, {"half_lat_deg", &m.half_lat_deg.d},
    {"area_cell", &m.area_cell.d}, {"area_lon", &m.area_lon.d}, {"area_lon_west", &m.area_lon_west.d},
    {"area_lon_east", &m.area_lon_east.d}, {"area_lon_north", &m.area_lon_north.d}, {"area_lon_south", &m.area_lon_south.d},
    {"area_lat", &m.area_lat.d}, {"area_lat_west", &m.area_lat_west.d}, {"area_lat_east", &m.area_lat_east.d},
    {"area_lat_north", &m.area_lat_north.d}, {"area_lat_south", &m.area_lat_south.d}, {"area_vtx", &m.area_vtx.d},
    {"area_subcell", &m.area_subcell.d}, {"de_lon", &m.de_lon.d}, {"le_lon", &m.le_lon.d}, {"de_lat", &m.de_lat.d},
    {"le_lat", &m.le_lat.d}, {"full_f", &m.full_f.d}, {"half_f", &m.half_f.d},
    {"full_tangent_wgt", &m.full_tangent_wgt.d}, {"half_tangent_wgt", &m.half_tangent_wgt.d},
    {"half_lon", &m.half_lon.d}, {"full_lon", &m.full_lon.d},
    {"hyai", &h->vc.hyai}, {"hybi", &h->vc.hybi}, {"full_lev", &m.full_lev.d}, {"half_lev", &m.half_lev.d},
    {"full_dlev", &m.full_dlev.d}, {"half_dlev", &m.half_dlev.d},
  };
  auto it = v.find(name);
  if (it == v.end()) return -1;
  if (buf) std::memcpy(buf, it->second->data(), it->second->size() * sizeof(double));
  return (int)it->second->size();
}

int gmh_filter_dims(const gmh_host* h, int which, int* ld_lon, int* ld_lat) {
  if (which < 0 || which > 2) return -1;
  *ld_lon = h->filt[which].ld_lon; *ld_lat = h->filt[which].ld_lat;
  return 0;
}
int gmh_get_filter(const gmh_host* h, int which, int* ngrid_lon, double* wgt_lon, int* ngrid_lat, double* wgt_lat) {
  if (which < 0 || which > 2) return -1;
  const Filter& f = h->filt[which];
  std::memcpy(ngrid_lon, f.ngrid_lon.data(), f.ngrid_lon.size() * sizeof(int));
  std::memcpy(ngrid_lat, f.ngrid_lat.data(), f.ngrid_lat.size() * sizeof(int));
  std::memcpy(wgt_lon, f.wgt_lon.data(), f.wgt_lon.size() * sizeof(double));
  std::memcpy(wgt_lat, f.wgt_lat.data(), f.wgt_lat.size() * sizeof(double));
  return 0;
}

int gmh_set_ic(const gmh_host* h, int test_case, double* u, double* v, double* pt_or_gz, double* phs, double* gzs,
               unsigned long long seed) {
  const Mesh& m = h->mesh;
  Lay L{m.hx, m.hy, m.nlon, m.nlat, m.nlev};
  switch (test_case) {
    case GMH_IC_ROSSBY_HAURWITZ: ic_rossby_haurwitz(h->c, m, L, u, v, pt_or_gz, gzs); return 0;
    case GMH_IC_MOUNTAIN_ZONAL_FLOW: ic_mountain_zonal_flow(h->c, m, L, u, v, pt_or_gz, gzs); return 0;
    case GMH_IC_BAROCLINIC_WAVE:
      ic_baroclinic_wave(h->c, m, h->vc, L, h->ptop, 1.0e5, u, v, pt_or_gz, phs, gzs); return 0;
    case GMH_IC_HELD_SUAREZ:
      ic_held_suarez(h->c, m, h->vc, L, h->ptop, 1.0e5, u, v, pt_or_gz, phs, gzs, seed, true); return 0;
    default: return -1;
  }
}

}  // extern "C"
