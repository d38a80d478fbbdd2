// Hydrostatic-core kernels (placeholder while the shallow-water path is validated).
#pragma once
#include "handle.cuh"
static int hydro_alloc(gmc_handle* h) { snprintf(h->err, sizeof(h->err), "hydrostatic core not built yet"); return 1; }
static int hydro_set_mesh(gmc_handle*, const gmc_mesh_vectors*) { return 0; }
static bool hydro_work_field(gmc_handle*, const std::string&, double**, int*) { return false; }
static void hydro_step(gmc_handle*, DState&, DState&, DState&, double) {}
static void hydro_post_step(gmc_handle*, int, double, int) {}
static void hydro_prepare_run(gmc_handle*, int) {}
static int hydro_check_flags(gmc_handle*) { return 0; }
static int hydro_set_tracers(gmc_handle* h, const double*) { snprintf(h->err, sizeof(h->err), "no tracers"); return 1; }
static int hydro_get_tracers(gmc_handle* h, double*) { snprintf(h->err, sizeof(h->err), "no tracers"); return 1; }
