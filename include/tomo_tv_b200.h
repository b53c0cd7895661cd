/*
 * tomo_tv_b200 -- C ABI of the B200-native tomo_TV iterative-reconstruction engine.
 *
 * This is the drop-in boundary for the reference's GPU path: one function per method the
 * reference binds on its pybind11 class `tomoengine`
 * (/root/reference/tomofusion/gpu/utils/tomoengine.cpp:487-534), with the same names, argument
 * meaning and host-side data layouts, so the reference's Python layer (TomoGPU,
 * tomofusion/gpu/reconstructor.py:12-215; pytvlib.initialize_algorithm/run, tomofusion/pytvlib.py:5-31)
 * can sit on top unchanged.  Plain pointers and sizes only; no torch / pybind / ASTRA types.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on error; ttv_last_error() gives the message
 *     (the reference aborts the process through ASTRA_ASSERT or ignores CUDA errors).
 *   - host arrays are float32, C-contiguous, in the reference's layouts:
 *       volume slice  [Ny][Nz]            (Matrix3D::getData/setData, container/Matrix3D.cpp:25-42)
 *       tilt series   [Nslice][Nproj*Nray], b[s][a*Nray + d]   (reconstructor.py:54-56, tomoengine.cpp:101)
 *   - inputs are copied; getters fill caller-provided buffers (ownership as in the reference).
 *   - calls are synchronous with respect to the host, like the reference's bound methods.
 *   - there is no CPU fallback: every entry point fails if no CUDA device is usable.
 */
#ifndef TOMO_TV_B200_H
#define TOMO_TV_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)   /* the library is built with -fvisibility=hidden */
#endif

typedef struct ttv_engine* ttv_handle;

const char* ttv_last_error(void);
int ttv_version(void);
int ttv_device_count(int* count);                 /* replaces tomofusion.device_count, tomofusion/__init__.py:10-18 */

/* tomoengine(Nslice, Nray) / tomoengine(Nslice, Nray, angles[rad])   tomoengine.cpp:34-84, :491-492.
 * device < 0: current device.  angles may be NULL with nangles == 0. */
int ttv_create(ttv_handle* out, int nslice, int nray, const float* angles_rad, int nangles, int device);
/* Slab constructor for tilt-axis sharding (replaces multigpuengine's OpenMP slice fan-out,
 * multigpuengine.cpp:113-228): this engine owns global slices [x_offset, x_offset+nslice_local). */
int ttv_create_slab(ttv_handle* out, int nslice_local, int nray, const float* angles_rad, int nangles,
                    int device, int x_offset, int nslice_global);
int ttv_destroy(ttv_handle h);

/* --- data in / out ------------------------------------------------------------------------ */
int ttv_set_tilt_series(ttv_handle h, const float* b, int rows, int cols);          /* tomoengine.cpp:101 */
int ttv_get_projections(ttv_handle h, float* out);                                  /* :456 */
int ttv_get_model_projections(ttv_handle h, float* out);                            /* :459 */
int ttv_initialize_initial_volume(ttv_handle h);                                    /* :95  */
int ttv_initialize_recon_copy(ttv_handle h);                                        /* :98  */
int ttv_set_original_volume(ttv_handle h, const float* slice, int s);               /* :104 */
int ttv_set_recon(ttv_handle h, const float* slice, int s);                         /* :106 */
int ttv_get_recon(ttv_handle h, float* out, int s);                                 /* :453 */
int ttv_update_projection_angles(ttv_handle h, const float* angles_rad, int nangles); /* :130-149 */
int ttv_create_projections(ttv_handle h);                                           /* :109-127 */
int ttv_restart_recon(ttv_handle h);                                                /* :462-468 */
int ttv_get_gpu_id(ttv_handle h, int* id);                                          /* :92 */
int ttv_set_gpu(ttv_handle h, int id);                                              /* :87-89 */
int ttv_poisson_noise(ttv_handle h, int Nc);                                        /* :471-484 */
/* bulk variants (additions): whole volume [Nslice][Ny][Nz] in one call */
int ttv_get_recon_all(ttv_handle h, float* out);
int ttv_set_recon_all(ttv_handle h, const float* in);
int ttv_set_original_volume_all(ttv_handle h, const float* in);

/* --- algorithms --------------------------------------------------------------------------- */
int ttv_initialize_SART(ttv_handle h, const char* order);                           /* :151-159 */
int ttv_SART(ttv_handle h, float beta, int niter);                                  /* :162-179 */
int ttv_initialize_SIRT(ttv_handle h);                                              /* :181-186 */
int ttv_SIRT(ttv_handle h, int niter);                                              /* :189-205 */
int ttv_initialize_FP(ttv_handle h);                                                /* :249-254 */
int ttv_initialize_BP(ttv_handle h);                                                /* :271-276 */
int ttv_initialize_poisson_ML(ttv_handle h);                                        /* :231-246 */
int ttv_poisson_ML(ttv_handle h, float lambda, float* cost);                        /* :293-315 */
int ttv_forward_projection(ttv_handle h);                                           /* :416-427 */
int ttv_initialize_fista(ttv_handle h);                                             /* :350-372 */
int ttv_get_lipschitz(ttv_handle h, float* L);                                      /* :375 */
int ttv_fista_momentum(ttv_handle h, float beta);                                   /* :381-384 */
int ttv_remove_momentum(ttv_handle h);                                              /* :378 */
int ttv_initialize_CGLS(ttv_handle h);                                              /* :207-211 */
int ttv_CGLS(ttv_handle h, int niter);                                              /* :214-229 */
int ttv_initialize_FBP(ttv_handle h, const char* filter);                           /* :317-327 */
int ttv_FBP(ttv_handle h, int apply_positivity);                                    /* :330-347 */

/* --- regularisers, elementwise, reductions ---------------------------------------------------- */
int ttv_tv_fgp(ttv_handle h, int ng, float lambda, float* tv_of_input);             /* :450 -> regularizers/tv_fgp.cu:192-281 */
int ttv_tv_gd(ttv_handle h, int ng, float dPOCS, float* tv_of_input);               /* :447 -> regularizers/tv_gd.cu:141-218 */
int ttv_tv(ttv_handle h, float* tv);                                                /* :441 */
int ttv_original_tv(ttv_handle h, float* tv);                                       /* :444 */
int ttv_soft_threshold(ttv_handle h, float lambda);                                 /* :436-438 */
int ttv_l1_norm(ttv_handle h, float* v);                                            /* :433 */
int ttv_copy_recon(ttv_handle h);                                                   /* :404 */
int ttv_matrix_2norm(ttv_handle h, float* v);                                       /* :407 */
int ttv_data_distance(ttv_handle h, float* v);                                      /* :410-413 */
int ttv_rmse(ttv_handle h, float* v);                                               /* :430 */

/* --- multi-GPU plumbing (one process per GPU; NCCL halo exchange + scalar all-reduce) --------
 * Replaces the unbuilt MPI variant's MPI_Isend/Recv + MPI_Allreduce
 * (tomofusion/gpu/utils/mpi_astra_ctvlib.cpp:360-386,160). */
int ttv_nccl_unique_id(void* id128);                       /* rank 0: 128-byte ncclUniqueId to broadcast */
int ttv_comm_init(ttv_handle h, const void* id128, int rank, int nranks);

/* Peer-memory halo hand-off for TV-FGP (optional, same node): every rank exports a 64-byte CUDA IPC
 * handle of its ghost-plane buffer, the handles are exchanged by the launcher (e.g. torch.distributed
 * all_gather) and each rank attaches its left / right neighbour's (NULL at the ends of the chain).
 * The sweep kernel then stores its boundary slices directly into the neighbours' HBM over NVLink and
 * publishes an epoch flag; no NCCL call is left on the per-sweep path. */
int ttv_p2p_export(ttv_handle h, void* handle64);
int ttv_p2p_attach(ttv_handle h, const void* left_handle64, const void* right_handle64);
/* same hand-off between engines living in ONE process (plain device pointers instead of IPC handles);
 * ttv_set_rank gives such an engine its position in the chain without an NCCL communicator -- the
 * caller then drives the slab engines from one host thread each (tests: two slabs on one GPU). */
int ttv_p2p_attach_local(ttv_handle h, ttv_handle left, ttv_handle right);
int ttv_set_rank(ttv_handle h, int rank, int nranks);

/* --- multi-modal data fusion: one entry point per bound method of the reference's pybind11 class
 * `multimodal` (/root/reference/tomofusion/chemistry/utils/multimodal.cpp:520-564), same names, argument
 * meaning and host layouts:
 *   HAADF tilt series   [Nslice][NprojH*Nray]                 bh[s][a*Nray + d]          (:156)
 *   chemical tilt series [Nslice][Nel*NprojC*Nray]            bChem[s][e*NprojC*Nray + a*Nray + d]   (:159, :342)
 *   reconstruction       recon[e][s][y][z] (Matrix4D.cpp:34-36), moved one [Ny][Nz] slice at a time (:495)
 *   sigma                [3][nnz] float32 rows / cols / vals  (:268-274, chemistry/reconstructor.py:150-153) */
typedef struct ttv_multimodal* ttv_mm_handle;
int ttv_mm_create(ttv_mm_handle* out, int nslice, int nray, int nelements, const float* haadf_angles_rad, int nhaadf,
                  const float* chem_angles_rad, int nchem, int device);                                /* :49-109 */
int ttv_mm_create_slab(ttv_mm_handle* out, int nslice_local, int nray, int nelements, const float* haadf_angles_rad, int nhaadf,
                       const float* chem_angles_rad, int nchem, int device, int x_offset, int nslice_global);
int ttv_mm_destroy(ttv_mm_handle h);
int ttv_mm_initialize_initial_volume(ttv_mm_handle h);                                           /* :120-121 */
int ttv_mm_set_haadf_tilt_series(ttv_mm_handle h, const float* bh, int rows, int cols);          /* :156 */
int ttv_mm_set_chem_tilt_series(ttv_mm_handle h, const float* bchem, int rows, int cols);        /* :159 */
int ttv_mm_set_measureHaadf(ttv_mm_handle h, int flag);                                          /* :127 */
int ttv_mm_set_measureChem(ttv_mm_handle h, int flag);                                           /* :133 */
int ttv_mm_measureHaadf(ttv_mm_handle h, int* flag);                                             /* :124 */
int ttv_mm_measureChem(ttv_mm_handle h, int* flag);                                              /* :130 */
int ttv_mm_set_gamma(ttv_mm_handle h, float gamma);                                              /* :145-153 */
int ttv_mm_gamma(ttv_mm_handle h, float* gamma);                                                 /* :142 */
int ttv_mm_set_recon(ttv_mm_handle h, const float* slice, int element, int s);                   /* :162 */
int ttv_mm_get_recon(ttv_mm_handle h, float* out, int element, int s);                           /* :495 */
int ttv_mm_set_original_volume(ttv_mm_handle h, const float* slice, int element, int s);         /* :165 */
int ttv_mm_get_gpu_id(ttv_mm_handle h, int* id);                                                 /* :117 */
int ttv_mm_set_gpu(ttv_mm_handle h, int id);                                                     /* :112-114 */
int ttv_mm_load_sigma(ttv_mm_handle h, const float* triplets, int nnz);                          /* :268-274 */
int ttv_mm_initialize_FP(ttv_mm_handle h);                                                       /* :168-171 */
int ttv_mm_initialize_BP(ttv_mm_handle h);                                                       /* :174-177 */
int ttv_mm_initialize_SIRT(ttv_mm_handle h);                                                     /* :331-336 */
int ttv_mm_initialize_SART(ttv_mm_handle h, const char* order);                                  /* :374-382 */
int ttv_mm_forward_projection(ttv_mm_handle h, const float* slice, float* out);                  /* :180-193 */
int ttv_mm_estimate_lipschitz(ttv_mm_handle h);                                                  /* :259-265 */
int ttv_mm_rescale_projections(ttv_mm_handle h);                                                 /* :312-328 */
int ttv_mm_rescale_tomograms(ttv_mm_handle h, float scale);                                      /* :307-309 */
int ttv_mm_poisson_ml(ttv_mm_handle h, float lambda_chem, float* cost);                          /* :277-304 */
int ttv_mm_data_fusion(ttv_mm_handle h, float lambda_haadf, float lambda_chem, int niter, const char* method,
                       float* cost_haadf, float* cost_chem);                                      /* :452-491 */
int ttv_mm_sirt_data_fusion(ttv_mm_handle h, float lambda_haadf, float lambda_chem, int niter, float* cost_haadf, float* cost_chem); /* :444-445 */
int ttv_mm_sart_data_fusion(ttv_mm_handle h, float lambda_haadf, float lambda_chem, float* cost_haadf, float* cost_chem);           /* :448-449 */
int ttv_mm_chemical_SIRT(ttv_mm_handle h, int niter);                                            /* :364-371 */
int ttv_mm_chemical_SART(ttv_mm_handle h, int niter);                                            /* :415-423 */
int ttv_mm_data_distance(ttv_mm_handle h, float* v);                                             /* :207-218 */
int ttv_mm_tv_gd(ttv_mm_handle h, int ng, float dPOCS, float* tv);                               /* :494 -> chemistry regularizers/tv_gd.cu */
int ttv_mm_tv_fgp_4D(ttv_mm_handle h, int ng, float lambda, float* tv);                          /* :497 -> chemistry regularizers/tv_fgp.cu:192-281 */
int ttv_mm_rmse(ttv_mm_handle h, float* per_element);                                            /* :503-506 */
int ttv_mm_get_haadf_projections(ttv_mm_handle h, float* out);                                   /* :513 */
int ttv_mm_get_chem_projections(ttv_mm_handle h, float* out);                                    /* :514 */
int ttv_mm_get_model_projections(ttv_mm_handle h, float* out);                                   /* :512 */
int ttv_mm_restart_recon(ttv_mm_handle h);                                                       /* :517 */
/* additions: bulk element volume [Nslice][Ny][Nz], launch counter, stream timer */
int ttv_mm_get_recon_all(ttv_mm_handle h, float* out, int element);
int ttv_mm_set_recon_all(ttv_mm_handle h, const float* in, int element);
int ttv_mm_launch_count(ttv_mm_handle h, long long* n);
int ttv_mm_timer_start(ttv_mm_handle h);
int ttv_mm_timer_stop(ttv_mm_handle h, float* ms);
int ttv_mm_comm_init(ttv_mm_handle h, const void* id128, int rank, int nranks);

/* --- measurement / test hooks (not part of the reference API) ------------------------------- */
int ttv_sync(ttv_handle h);
/* device-side __cosf/__sinf of the per-angle FP angle, as the projector evaluates them */
int ttv_get_trig(ttv_handle h, float* cosv, float* sinv);
/* R = inv(FP(1)) [Nproj*Nray], C = inv(BP(1)) [Ny*Nz]   (A/cuda/2d/sirt.cu:112-147) */
int ttv_get_sirt_weights(ttv_handle h, float* R, float* C);
/* back-project the model projections g into recon (vol = BP(g)); operator-level parity tests */
int ttv_back_projection(ttv_handle h);
int ttv_set_model_projections(ttv_handle h, const float* g, int rows, int cols);
/* time `reps` back-to-back launches of one named kernel group with CUDA events on the engine's
 * stream; *ms = average per launch.  names: fp, bp, sirt, fgp_sweep, fgp_first, fgp_final, fgp,
 * tv, momentum, tvgd_grad, tvgd_update, sqdist, fista_iter, kl, sart. */
int ttv_time_kernel(ttv_handle h, const char* name, int reps, int arg_i, float arg_f, float* ms);
/* per-kernel-group CUDA-event timing on the engine's stream: enable (resets the counters), run, then
 * read total milliseconds and launch count per group.  groups: fp, bp, fgp_first, fgp_sweep, fgp_final,
 * tv, momentum, tvgd_grad, tvgd_update, reduce, elementwise, sart_fp, sart_bp, io, halo, fgp2_first, fgp2_sweep,
 * fgp2_final (two-sweep launches). */
int ttv_profile(ttv_handle h, int enable);
int ttv_profile_get(ttv_handle h, const char* group, double* total_ms, long long* count);
/* CUDA events on the engine's stream around an arbitrary sequence of calls */
int ttv_timer_start(ttv_handle h);
int ttv_timer_stop(ttv_handle h, float* ms);
/* implementation switches for A/B tests: "fgp_tiled" (1 = TMA-tiled TV-FGP sweep, 0 = first-generation kernel);
 * "fgp_fuse" (1 = two FGP sweeps per launch on a single-volume engine, ttv_fgp2.cuh; 0 = one launch per sweep);
 * "fp_variant" / "bp_variant" (0 = by slab thickness, 1 = classic fp_kernel / bp_kernel, 2 = table-driven fpt_kernel /
 * bpt_kernel); "fp_xb" (x-blocks per CTA of fpt_kernel: 0 = by slab thickness and wave count, 1, 2, 3 = both sizes in one launch,
 * fptm_kernel, with "fp_mix_n2" leading two-x-block items, -1 = half of them); "bp_xb" (x-blocks per CTA of bpt_kernel: 0 = by slab thickness, 2, 8); "sart_cluster" (1 = slice-resident cluster
 * kernel of ttv_sartc.cuh where at least two slices fit in a cluster's shared memory, 2 = also with one slice per cluster, 0 = one
 * launch pair per angle); "tvgd_march" (1 = register-marching TV-GD gradient kernel, 0 = first-generation kernel).  Every value
 * selects a kernel the GPU tests validate (projector, TV-FGP and TV-GD variants are bit-identical to each other; the two SART
 * paths agree to float32 rounding). */
int ttv_set_option(ttv_handle h, const char* name, int value);
/* self-test of the inlined 1/sqrt used by the TV-FGP projection step: number of n pseudo-random arguments > 1 whose result
 * differs from IEEE div(1, sqrt(x)) (the reference's expression, tv_fgp.cu:106); must be 0 */
int ttv_selftest_invsqrt(ttv_handle h, unsigned long long n, unsigned seed, unsigned long long* mismatches);
/* number of kernel launches issued by this engine so far */
int ttv_launch_count(ttv_handle h, long long* n);
/* algorithmic geometry facts for the roofline: bytes of one volume (V) and one sinogram (S) */
int ttv_sizes(ttv_handle h, long long* V, long long* S, long long* samples_per_projection);

/* TV value of a DEVICE array without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that
 * forbid it).  Same arithmetic as ttv_tv (tv_3D_kernel, tv_gd.cu:27-47: sum of sqrt(1e-6 + dx^2 + dy^2 + dz^2), periodic
 * forward differences) on the caller's own layout:
 *   d_vol   device pointer, float32 [nx][ny][nz], C-contiguous (nx slices of [ny][nz], the layout of ttv_get_recon_all)
 *   d_tv    device pointer to ONE float32: receives the TV value (stream-ordered; the call does not synchronise)
 *   stream  a cudaStream_t, or NULL for the default stream; the current device must be the one that owns the pointers
 * No engine handle, no cudaMalloc / cudaFree or allocating library call: the reduction scratch (one partial sum per 8 rows
 * + a ticket) is a static __device__ array of TTV_NOALLOC_MAX_ROWS / 8 doubles + 256 bytes (1 MiB + 256 B) that every GPU
 * loading the library holds until exit.  Limit: nx * ny <= TTV_NOALLOC_MAX_ROWS (e.g. 1024 x 1024 x any nz); larger sizes are
 * rejected with an error.  All calls share the scratch array: calls that OVERLAP on the GPU (different streams or host
 * threads) race on it and must be serialised by the caller.  The sum is taken in a fixed order: repeated calls on the same
 * input give the same bits. */
#define TTV_NOALLOC_MAX_ROWS 1048576
int ttv_tv_value_noalloc(const float* d_vol, int nx, int ny, int nz, float* d_tv, void* stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif
