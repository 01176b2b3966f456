/* pace_b200 — C ABI of the B200-native FV3 dycore hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch / C++ types.
 * The host-side mirror of the reference classes (pace_b200/engine.py, pace_b200/dropin.py) binds these with ctypes; the
 * reference-side binding a maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - every call returns 0 on success, non-zero on failure; pace_b200_last_error() gives the text.
 *   - all arrays are DEVICE pointers to fp64, owned by the caller (PyTorch), laid out
 *       3-D: f[t][k][j][i]  (t = subdomain on this GPU, k in [0,npz], j in [0,ny+7), i in [0,sj))
 *       2-D: f[t][j][i],  K-only: f[k]
 *     with the reference's index convention: 3-cell halo, origin (3,3,0) (NDSL sizer.py:131-154).
 *   - fields are registered once by name (the reference's own attribute names, e.g. "delp",
 *     "ad.gz" for AcousticDynamics._gz, "grid.rarea" for GridData.rarea); module calls then take
 *     only scalars, like the reference's __call__ signatures minus the arrays.
 *   - calls are asynchronous on the stream given to pace_b200_set_stream (default stream 0).
 */
#ifndef PACE_B200_H
#define PACE_B200_H

#ifdef __cplusplus
extern "C" {
#endif

/* Subset of pyFV3.DynamicalCoreConfig (pyFV3/_config.py:143-281) read on the hot path, plus the
 * subdomain geometry of NDSL SubtileGridSizer (sizer.py:20-154). */
typedef struct pace_b200_config {
  int npx, npy, npz;
  int nx, ny;          /* compute cells per subdomain */
  int n_tiles_local;   /* subdomains resident on this GPU */
  int n_split, k_split;
  int hord_dp, hord_mt, hord_tm, hord_tr, hord_vt;
  int kord_mt, kord_tm, kord_tr, kord_wz;
  int nord, n_sponge, nf_omega;
  int grid_type;
  int do_sat_adj, do_vort_damp, rf_fast, hydrostatic, convert_ke, fill, adiabatic, do_qa;
  int use_old_omega, do_f3d, inline_q, nwat;
  double dt_atmos;
  double a_imp, beta, p_fac, d_con, d2_bg, d2_bg_k1, d2_bg_k2, d4_bg, dddmp, delt_max, ke_bg;
  double rf_cutoff, tau, vtdm4, d_ext, consv_te;
  double ptop, da_min, da_min_c;
  int ks;
} pace_b200_config;

/* Row stride (elements) the library expects for a subdomain of nx compute cells. */
int pace_b200_row_stride(int nx);

/* Create a dycore instance for `cfg->n_tiles_local` subdomains on the current CUDA device.
 * edge_flags[t]: bit0 west, bit1 east, bit2 south, bit3 north — whether subdomain t touches that
 * edge of its cube tile (GridIndexing.west_edge..., NDSL dsl/stencil.py:560-620). */
int pace_b200_create(const pace_b200_config* cfg, const unsigned char* edge_flags, void** handle);
int pace_b200_destroy(void* handle);
const char* pace_b200_last_error(void* handle);
int pace_b200_set_stream(void* handle, void* cuda_stream);
/* Register a caller-owned device array under the reference's attribute name. */
int pace_b200_set_field(void* handle, const char* name, double* device_ptr);

/* Measurement hooks (bench.py): number of kernels launched by this library so far; CUDA-event timing of the
 * kernels that carry a label (enable, run, then report "label count total_ms" lines and clear). */
long long pace_b200_launch_count(void);
int pace_b200_profile_enable(int on);
int pace_b200_profile_report(char* buf, int n);

/* Set a small per-level (K) column from HOST memory, e.g. the column namelist of d_sw
 * (pyFV3/stencils/d_sw.py:647-691: "col.nord", "col.nord_v", "col.damp_vt", ...) or the delnflux damping
 * columns (delnflux.py:22-38).  The library keeps a host copy and a device copy under `name`. */
int pace_b200_set_column(void* handle, const char* name, const double* host_values, int n);

/* NonhydrostaticVerticalSolverCGrid.__call__ (pyFV3/stencils/riem_solver_c.py:200-278).
 * in: csw.delpc csw.ptc ad.cappa q_con omga(w3) ad.ws3 phis(hs) ad.gz   out: ad.gz ad.pkc(pef) */
int pace_b200_riem_solver_c(void* handle, double dt2);

/* CGridShallowWaterDynamics.__call__ (pyFV3/stencils/c_sw.py:644-807) incl. DGrid2AGrid2CGridVectors
 * (d2a2c_vect.py:588-699).  in: delp pt u v w   out: uc vc ua va ad.ut ad.vt ad.divgd omga csw.delpc
 * csw.ptc (and the halo corners of delp, pt, w, as in the reference) */
int pace_b200_c_sw(void* handle, double dt2);

/* UpdateGeopotentialHeightOnCGrid.__call__ (pyFV3/stencils/updatedzc.py:185-226).
 * in: ad.zs ad.ut ad.vt ad.gz grid.area grid.dp_ref   out: ad.gz ad.ws3 */
int pace_b200_update_dz_c(void* handle, double dt2);

/* No counterpart in the reference.  Same computation, inputs, outputs and results as pace_b200_update_dz_c, for
 * callers that cannot let a library allocate device memory: a host model whose own pool took the whole GPU at
 * start-up, or a time-critical loop that cannot afford the device-wide synchronisation cudaMalloc implies.  Its one
 * work array (a copy of ad.gz) is a fixed array of PACE_B200_STATIC_SCRATCH_DOUBLES doubles inside the library image
 * (every GPU that loads the library holds those 16 MiB until the process exits); no allocation routine is called.
 * Returns non-zero, with a message in pace_b200_last_error, when the handle's fields (subdomains x (npz+1) x (ny+7)
 * x row stride doubles) exceed that size.  The first call on each device resolves the array's device address: make
 * it before starting any stream capture on that device.  All handles share the array and it keeps the last call's
 * contents (each call overwrites everything it reads, nothing needs clearing); calls that can overlap on one GPU
 * race on it, so order them on one stream or with events. */
#define PACE_B200_STATIC_SCRATCH_DOUBLES 2097152
int pace_b200_update_dz_c_static(void* handle, double dt2);

/* p_grad_c_stencil (pyFV3/stencils/dyn_core.py:122-173).  in: csw.delpc ad.pkc ad.gz   inout: uc vc */
int pace_b200_p_grad_c(void* handle, double dt2);

/* Pointwise stencils of DynamicalCore._compute: fv_setup (pyFV3/stencils/moist_cv.py:201-234),
 * pt_to_potential_density_pt and omega_from_w (fv_dynamics.py:33-58), copy_defn on npz levels. */
int pace_b200_fv_setup(void* handle);
int pace_b200_pt_to_potential_density_pt(void* handle);
int pace_b200_omega_from_w(void* handle);
int pace_b200_copy_levels(void* handle, const char* src, const char* dst);

/* TracerAdvection.__call__ (pyFV3/stencils/tracer_2d_1l.py:289-417) in the three pieces between which the
 * reference updates the tracer halos: flux_compute + divide_fluxes_by_n_substeps; one sub-cycle
 * (apply_mass_flux, then fvtp2d + apply_tracer_flux per tracer); swap_dp.
 * inout: the nq tracers (qvapor, qliquid, qrain, qice, qsnow, qgraupel, qo3mr, qsgs_tke), dyn.dp1, cxd cyd mfxd mfyd */
int pace_b200_tracer_flux_prep(void* handle, int n_split);
int pace_b200_tracer_substep(void* handle, int nq);
int pace_b200_tracer_swap_dp(void* handle);

/* Test hook: the kernels' exp / log / pow (csrc/libm_port.cuh, bit-identical to the oracle's host libm)
 * evaluated elementwise on device arrays.  op 0: exp(x), 1: log(x), 2: pow(x, y). */
int pace_b200_math_eval(void* handle, int op, int n, const double* x, const double* y, double* out);

/* MapSingle.__call__ (pyFV3/stencils/map_single.py:155-208; Fortran map_single / map_scalar / map1_ppm):
 * conservative PPM remap of q from source interfaces pe1 to target interfaces pe2, per column.
 * iv: -2 (w), -1 (winds), 0 (positive-definite tracers), 1 (others); kord <= 10; qs: optional bottom value (iv -2). */
int pace_b200_map_single(void* handle, const char* q, const char* pe1, const char* pe2, const char* qs, int iv,
                         int kord, int x_interface, int y_interface, double qmin);

/* The same MapSingle.__call__ in the reference's single-precision mode (PACE_FLOAT_PRECISION=32: Float = np.float32,
 * NDSL/ndsl/dsl/typing.py:25-45; every field, literal and intermediate of map_single.py / remap_profile.py is then
 * float32).  The handle's registry holds fp64 fields, so the fp32 fields are passed as raw device pointers; they use
 * the handle's layout ([t][k][j][i], row stride pace_b200_row_stride(nx) ELEMENTS, npz + 1 levels).  qs may be NULL. */
int pace_b200_map_single_f32(void* handle, float* q, const float* pe1, const float* pe2, const float* qs, int iv,
                             int kord, int x_interface, int y_interface, float qmin);

/* LagrangianToEulerian.__call__ (pyFV3/stencils/remapping.py:522-731) in two pieces around the saturation
 * adjustment: remap_main = init_pe .. copy_from_below (13 MapSingle + fillz); remap_finish = moist_pt_last_step
 * (last_step) or pt /= pkz. */
int pace_b200_remap_main(void* handle, int nq);
int pace_b200_remap_finish(void* handle, int last_step);

/* SatAdjustConfig (pyFV3/_config.py:15-66) as read by SatAdjust3d. */
typedef struct pace_b200_satadj_config {
  int rad_snow, rad_rain, rad_graupel, tintqs, icloud_f, do_qa;
  double sat_adj0, ql_gen, qs_mlt, ql0_max, t_sub, qi_gen, qi_lim, qi0_max, dw_ocean, dw_land, cld_min;
  double tau_i2s, tau_v2l, tau_r2g, tau_l2r, tau_l2v, tau_imlt, tau_smlt;
} pace_b200_satadj_config;

/* SatAdjust3d.__call__ (pyFV3/stencils/saturation_adjustment.py:982-1108) on levels [kmp, npz).
 * inout: qvapor qliquid qice qrain qsnow qgraupel pt ad.cappa   out: qcld q_con pkz   in: delp delz phis grid.area_64 */
int pace_b200_sat_adjust(void* handle, const pace_b200_satadj_config* cfg, double mdt, int last_step, int kmp);

/* AdjustNegativeTracerMixingRatio.__call__ (pyFV3/stencils/neg_adj3.py:385-431).
 * inout: qvapor qliquid qrain qsnow qice qgraupel qcld pt   in: delp */
int pace_b200_neg_adj3(void* handle);
/* CubedToLatLon ord4_transform (NDSL/ndsl/stencils/c2l_ord.py:110-157); the u/v halo update that precedes it
 * in CubedToLatLon.__call__ is issued by the caller.  in: u v   out: ua va */
int pace_b200_c2l_ord4(void* handle);

/* One region of a halo update: the destination box (array indices) of subdomain dst_t is filled from
 * subdomain src_t at index (aii*i + aij*j + bi, aji*i + ajj*j + bj); comp/src_comp select the x (0) or y (1)
 * array of a vector pair, sign its sign.  Replaces Boundary + rotate_*_data + the pack/unpack index arrays of
 * NDSL (comm/boundary.py:10-120, halo/rotate.py:4-50, halo/data_transformer.py:54-99). */
typedef struct pace_b200_halo_op {
  int dst_t, src_t, comp, src_comp;
  double sign;
  int i0, j0, ni, nj;
  int aii, aij, bi, aji, ajj, bj;
} pace_b200_halo_op;

/* Upload a plan (list of ops between subdomains of this GPU) under plan_id in [0,256). */
int pace_b200_halo_plan_set(void* handle, int plan_id, const void* ops, int n_ops);
/* HaloUpdater.update() for up to 8 fields (or vector pairs when y_names != NULL) sharing one plan
 * (NDSL/ndsl/halo/updater.py:150-303), levels [0, nk). */
int pace_b200_halo_update(void* handle, int plan_id, const char* const* x_names, const char* const* y_names,
                          int n, int nk);
/* The two halves of HaloUpdater.start()/wait() for neighbours that live on another GPU (NDSL/ndsl/halo/updater.py:217-303
 * pack + Isend / Irecv + unpack; data_transformer.py:404-560 with the rotation applied while packing).  A pack plan
 * holds ops whose source subdomain is local and whose dst_t field is the op's cell offset in the message; an unpack
 * plan holds ops whose destination is local and whose src_t field is that offset.  Message element of (op, field q,
 * level k, cell c) = (offset * n + q * cells) * nk + k * cells + c, doubles; both sides list the ops in the same order.
 * `message` is a device buffer the caller transfers (NCCL send/recv in pace_b200/distributed.py). */
int pace_b200_halo_pack(void* handle, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk,
                        void* message);
int pace_b200_halo_unpack(void* handle, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk,
                          const void* message);

/* FiniteVolumeFluxPrep.__call__ (pyFV3/stencils/fxadv.py:590-691).
 * in: uc vc   out: ad.crx ad.cry ad.xfx ad.yfx dsw.uc_contra dsw.vc_contra */
int pace_b200_fv_prep(void* handle, double dt);

/* DGridShallowWaterLagrangianDynamics.__call__ (pyFV3/stencils/d_sw.py:997-1303).
 * inout: delp pt u v w uc vc q_con mfxd mfyd cxd cyd ad.heat_source diss_estd ad.divgd ad.vt(delpc)
 * in: ua va ad.zh   out: ad.crx ad.cry ad.xfx ad.yfx */
int pace_b200_d_sw(void* handle, double dt);

/* NonhydrostaticVerticalSolver.__call__ (pyFV3/stencils/riem_solver3.py:225-339).
 * in: delp ad.cappa q_con pt ad.zs ad.wsd  inout: pe ad.zh w delz  out: ad.pkc ad.pk3 (+ pk peln if last_call) */
int pace_b200_riem_solver3(void* handle, double dt, int last_call);

/* UpdateHeightOnDGrid.__call__ (pyFV3/stencils/updatedzd.py:316-389).
 * in: ad.zs ad.crx ad.cry ad.xfx ad.yfx  inout: ad.zh  out: ad.wsd */
int pace_b200_update_dz_d(void* handle, double dt);

/* Helper stencils of AcousticDynamics (pyFV3/stencils/dyn_core.py:53-119, 603-607). */
int pace_b200_zero_data(void* handle, int first_timestep);
int pace_b200_gz_from_zs_delz(void* handle);
int pace_b200_pem_from_delp(void* handle);
int pace_b200_copy_full(void* handle, const char* src, const char* dst);
int pace_b200_compute_geopotential(void* handle);

/* edge_pe (pyFV3/stencils/pe_halo.py:6-34) and PK3Halo.__call__ (pk3_halo.py:11-73). */
int pace_b200_edge_pe(void* handle);
int pace_b200_pk3_halo(void* handle);

/* NonHydrostaticPressureGradient.__call__ (pyFV3/stencils/nh_p_grad.py:201-265).
 * inout: u v ad.pkc(pp) ad.gz ad.pk3   in: delp */
int pace_b200_nh_p_grad(void* handle, double dt);

/* RayleighDamping.__call__ (pyFV3/stencils/ray_fast.py:49-206).  inout: u v w */
int pace_b200_ray_fast(void* handle, double dt);

/* HyperdiffusionDamping.__call__ (pyFV3/stencils/del2cubed.py:163-202) on the registered field q. */
int pace_b200_del2cubed(void* handle, const char* q, double cd, int nmax);

/* apply_diffusive_heating (pyFV3/stencils/temperature_adjust.py:8-42) on levels [0, nk). */
int pace_b200_apply_diffusive_heating(void* handle, double delt_time_factor, int nk);

/* FiniteVolumeTransport.__call__ (pyFV3/stencils/fvtp2d.py:251-368) on registered fields; optional
 * arguments are "" (or NULL).  nord_col/damp select the DelnFlux damping (fvtp2d.py:168-180). */
int pace_b200_fvtp2d(void* handle, const char* q, const char* crx, const char* cry, const char* xfx,
                     const char* yfx, const char* fx, const char* fy, const char* x_mass_flux,
                     const char* y_mass_flux, const char* mass, int hord, const char* nord_col,
                     const char* damp, int nk);

/* DelnFluxNoSG.__call__ (pyFV3/stencils/delnflux.py:1243-1278). */
int pace_b200_delnflux_nosg(void* handle, const char* q, const char* fx2, const char* fy2, const char* damp,
                            const char* d2, const char* nord_col, int nk);

/* AGrid2BGridFourthOrder.__call__ (pyFV3/stencils/a2b_ord4.py:715-808) on levels [k0, k0+nk). */
int pace_b200_a2b_ord4(void* handle, const char* qin, const char* qout, int k0, int nk, int replace);

/* DivergenceDamping.__call__ (pyFV3/stencils/divergence_damping.py:549-713). */
int pace_b200_divergence_damping(void* handle, const char* vort_b, const char* delpc, const char* ke,
                                 const char* vort_a, double dt);

/* ---- Whole-phase entry points -------------------------------------------------------------------------------
 * The call order of the reference's Python driver classes, so that a caller in any language (the GEOS wrapper
 * precedent, pyFV3/wrappers/geos_wrapper.py:224-232) steps the core without re-stating it.  They use the module
 * calls above on the handle's stream and need, registered beforehand: every DycoreState field and "ad." / "csw." /
 * "dyn." temporary (INTEGRATION.md lists them), the grid, the K-columns, and halo plans under the ids
 *   0 cell-centred 3-cell halo, 1 cell-centred 2-cell, 2 corner (x- and y-interface) 3-cell, 3 D-grid vector pair,
 *   4 C-grid vector pair, 5 interface synchronisation of (u, v)   (pace_b200.halo.HaloPlanner builds all six). */

/* HaloUpdater.update() when not every neighbour subdomain lives on this GPU (NDSL/ndsl/halo/updater.py:150-303):
 * the phase calls hand the whole update to this callback (it packs with pace_b200_halo_pack, moves the messages and
 * unpacks, all ordered on the handle's stream) instead of running the local gather.  Return 0 on success. */
typedef int (*pace_b200_halo_callback)(void* user, int plan_id, const char* const* x_names, const char* const* y_names,
                                       int n, int nk);
int pace_b200_set_halo_callback(void* handle, pace_b200_halo_callback cb, void* user);

/* Halo exchange with subdomains on other GPUs of the node WITHOUT leaving the stream (replaces the Isend / Irecv of
 * NDSL/ndsl/halo/updater.py:217-303 and the mpi4py calls behind comm/communicator.py): every rank owns an "arena" of
 * 2 * half_doubles doubles followed by n_ranks 8-byte flag slots, zero-initialised and mapped into every peer process
 * (CUDA IPC / torch symmetric memory); arena_ptrs[r] is rank r's arena as mapped HERE.  For a plan id (0..15) the caller
 * uploads, per peer GPU, a pack plan whose message offsets follow that peer's receive layout, and one unpack plan
 * (pace_b200_halo_plan_set, ids >= 16); max_recv_cells = the largest number of halo cells any rank receives in one
 * update of the plan (the same number on every rank: updates that share a wait point — (u, v) with w, zh with pkc —
 * travel as one exchange, the second message behind the largest first one).  The phase calls then run pack-into-peer-arena, local gather, a flag barrier
 * across all ranks and unpack as kernels on the handle's stream; pace_b200_halo_peer_update does one such update on
 * its own.  Every rank must issue the same sequence of updates.  A barrier that waits 30 s for a peer gives up and the
 * next call reports it. */
int pace_b200_halo_peer_arena(void* handle, int my_rank, int n_ranks, void* const* arena_ptrs, long long half_doubles);
int pace_b200_halo_peer_plan(void* handle, int plan_id, int n_peers, const int* peer_ranks, const int* pack_plan_ids,
                             int unpack_plan_id, long long max_recv_cells);
int pace_b200_halo_peer_update(void* handle, int plan_id, const char* const* x_names, const char* const* y_names, int n,
                               int nk);
/* Waits for the stream's work, then: number of barriers executed, and the sequence number of a barrier that gave up
 * waiting for a peer (0 = none). */
int pace_b200_halo_peer_status(void* handle, unsigned long long* barriers_done, unsigned long long* timed_out);

/* The reference's checkpointer hooks (dyn_core.py:641-701, 787-893; fv_dynamics.py:342-443, 457-606): the callback is
 * called with the savepoint name ("FVDynamics-In", "C_SW-In", "C_SW-Out", "D_SW-In", "D_SW-Out", "Tracer2D1L-In",
 * "Tracer2D1L-Out", "Remapping-In", "Remapping-Out", "FVDynamics-Out") at the same points of the sequence.  The work
 * queued so far is NOT synchronised: a callback that reads fields synchronises the handle's stream first. */
typedef int (*pace_b200_stage_callback)(void* user, const char* savepoint);
int pace_b200_set_stage_callback(void* handle, pace_b200_stage_callback cb, void* user);

/* SatAdjustConfig used by pace_b200_remap / pace_b200_step when do_sat_adj is set. */
int pace_b200_set_satadj_config(void* handle, const pace_b200_satadj_config* cfg);

/* AcousticDynamics.__call__(state, timestep, n_map) (pyFV3/stencils/dyn_core.py:714-1013): n_split acoustic
 * substeps of timestep / n_split seconds incl. every halo update, then the heat-source diffusion. */
int pace_b200_acoustic(void* handle, double timestep, int n_map);
/* TracerAdvection.__call__ (pyFV3/stencils/tracer_2d_1l.py:289-417) for the eight GFS tracers. */
int pace_b200_tracer_advection(void* handle);
/* LagrangianToEulerian.__call__ (pyFV3/stencils/remapping.py:522-731) incl. SatAdjust3d when do_sat_adj. */
int pace_b200_remap(void* handle, int last_step, double mdt);
/* DynamicalCore.step_dynamics (pyFV3/stencils/fv_dynamics.py:445-655): one dycore step of dt_atmos. */
int pace_b200_step(void* handle);

/* ---- dycore <-> physics coupling of the driver's time loop (driver/pace/driver/driver.py:645-661) -------------------
 * Fields, registered like all others: "phy.<name>" = PhysicsState.<name> (the 16 copied fields qvapor ... omga, prsi and
 * physics_updated_{ua,va,pt,specific_humidity,qliquid,qrain,qsnow,qice,qgraupel}), "tend.u_dt / v_dt / pt_dt" =
 * TendencyState, "dgrid.<name>" = DriverGridData (vlon1-3, vlat1-3, edge_vect_w/e/s/n, es1_1-3, ew2_1-3; the two
 * edge vectors that are 1-D along i in the reference as 2-D fields constant along j).  The physics parameterisations
 * are the caller's: it fills phy.prsi and phy.physics_updated_* between the two calls. */
/* DycoreToPhysics.__call__ with do_dry_convective_adjust = false (pySHiELD/update/update_atmos_state.py:184-231). */
int pace_b200_dycore_to_physics(void* handle);
/* UpdateAtmosphereState.__call__(dycore_state, phy_state, u_dt, v_dt, pt_dt, dt) (update_atmos_state.py:296-349) incl.
 * ApplyPhysicsToDycore (fv_update_phys.py:152-195), AGrid2DGridPhysics (update_dwind_phys.py:490-701), the halo updates
 * of the tendencies and of the winds, and CubedToLatLon.  dycore_only / apply_tendencies: the constructor arguments of
 * the reference class.  origin_i / origin_j (one int per subdomain of the handle): subtile_index * subtile_width, what
 * the reference adds to a local index to get its position on the tile (local_to_global_1d); npx, npy: namelist. */
int pace_b200_update_atmos_state(void* handle, double dt, int dycore_only, int apply_tendencies, const int* origin_i,
                                 const int* origin_j, int npx, int npy);

/* ---- diagnostics egress (driver/pace/driver/diagnostics.py:142-249) ------------------------------------------------
 * What MonitorDiagnostics.store hands its monitor, produced on the device in the reference's own order.
 * pace_b200_egress_pack: `quantity.view[:]` of a field as a dense [subdomain][i][j][k] array (k fastest) at the device
 * address dst.  kind 0: 3-D field, levels [0, nk); kind 1: 2-D field; kind 2: level `level` of a 3-D field as a 2-D array
 * (ZSelect.select_data, diagnostics.py:44-64).  x_interface / y_interface: the field has nx+1 / ny+1 points.
 * pace_b200_column_integral: _compute_column_integral (diagnostics.py:226-249) into the registered 2-D field out2d,
 * bit-identical to RGRAV * np.sum(q * delp, axis=2). */
int pace_b200_egress_pack(void* handle, const char* name, int kind, int level, int x_interface, int y_interface, int nk,
                          double* dst);
int pace_b200_column_integral(void* handle, const char* q, const char* delp, const char* out2d);

#ifdef __cplusplus
}
#endif
#endif
