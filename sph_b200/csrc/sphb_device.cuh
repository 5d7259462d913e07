// Device-side code of the SPH step engine: grid build (cell keys, stable counting sort, reorder),
// density/pressure, fused force+integrate+collide, AoS<->SoA conversion.  sm_100a only.
//
// Reference restated (not ported: the reference is a scalar all-pairs CPU loop):
//   SPH/SphUtils.cpp:8-47   update_cells      -> k_bbox, k_grid_setup, k_key_hist, k_scan_*, k_scatter, k_rank_gather
//   SPH/SphUtils.cpp:55-95  W, gradW, lapW    -> w_shape(), grad_shape(), k_eval_kernels
//   SPH/SphUtils.cpp:100-149 update_density   -> k_density
//   SPH/SphUtils.cpp:157-239,295-304 forces   -> k_force<...>
//   SPH/SphUtils.cpp:321-335 update_posvel    -> integrate()
//   SPH/SphUtils.cpp:343-372 collision (walls)-> collide_walls()
#pragma once
#include "sphb_common.cuh"
#include "sphb_grid.cuh"
#include "sphb_density.cuh"
#include "sphb_force.cuh"
#include "sphb_io.cuh"
#include "sphb_spheres.cuh"
#include "sphb_render.cuh"
#include "sphb_multigpu.cuh"
#include "sphb_parity.cuh"
