/*
 * include/dis_b200.h — C ABI of the B200-native Dense Inverse Search engine.
 *
 * This is the drop-in boundary for the reference's hot path.  The reference has no FFI: its
 * boundary is the C++ class API.  Each entry point below names the reference
 * interface it replaces (paths relative to /root/reference).  Plain pointers and sizes only; no
 * C++/torch types; every function returns DIS_OK (0) or a negative dis_status and records a
 * message retrievable with dis_last_error().  No exceptions cross this boundary.
 *
 * There is NO CPU fallback: every compute entry point runs hand-written sm_100a kernels and
 * fails with DIS_ERR_CUDA when no usable device is present.
 *
 * Threading: a handle is used by one host thread at a time; it owns one CUDA stream and all the
 * device memory it allocates.
 */
#ifndef DIS_B200_H
#define DIS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define DIS_API __attribute__((visibility("default")))
#else
#define DIS_API
#endif

typedef enum dis_status {
  DIS_OK = 0,
  DIS_ERR_PARAM = -1,    /* invalid parameter / size (the reference would read out of bounds) */
  DIS_ERR_CUDA = -2,     /* CUDA runtime error, or no sm_100 device */
  DIS_ERR_STATE = -3,    /* call sequence error (e.g. results queried before a run) */
  DIS_ERR_ALLOC = -4
} dis_status;

/* The parameters of src/main.cpp:63-72 (same names, same defaults via dis_default_params):
 * max_iter=1000, patch_size=8, coarsest_scale=3, finest_scale=0, patch_overlap=0.7f,
 * patch_norm=1.  draw_grid is a GUI path and is not part of this ABI. */
typedef struct dis_params {
  int max_iter;          /* `iterations`: a patch runs max_iter+1 passes unless it is reset (src/patch.cpp:199) */
  int patch_size;        /* even, >= 4 */
  int coarsest_scale;
  int finest_scale;
  float patch_overlap;   /* steps = max(1, floor(patch_size*(1-overlap))) in fp32 (src/optical_flow.cpp:38) */
  int patch_norm;        /* mean-normalise the sampled patch (src/patch.cpp:264) */
  /* Extension, off by default (0): stop a patch once |delta_u| < conv_eps.  The reference has no
   * convergence criterion; conv_eps > 0 is a reported deviation, never the default. */
  float conv_eps;
  /* fix_parameters fields the reference derives but its inner classes take explicitly
   * (include/optical_flow.hpp:33-35); 0 = derive them as src/optical_flow.cpp:34,38 does. */
  int steps;             /* distance between patch centres; 0: max(1, floor(patch_size*(1-overlap))) */
  float outlier_thresh;  /* displacement (px) beyond which a patch is reset; 0: patch_size/2 */
  int reserved[3];
} dis_params;

typedef struct dis_engine dis_engine;   /* opaque handle */

/* Geometry of one level (src/optical_flow.cpp:49-59, src/patch_grid.cpp:20-25). */
typedef struct dis_level_info {
  int scale;             /* absolute scale */
  int width, height;     /* level size */
  int tmp_w, tmp_h;      /* padded plane size */
  int npw, nph;          /* patch grid */
  int offw, offh;        /* grid offsets */
  int steps;
  int num_patches;
} dis_level_info;

/* Counters of the last run (summed over pairs and levels). */
typedef struct dis_stats {
  unsigned long long patches;           /* patches processed */
  unsigned long long patch_iterations;  /* passes of the reference loop accounted for (src/patch.cpp:165-202) */
  unsigned long long sampled_iterations;/* passes for which the 4-tap resample + reductions actually ran */
  unsigned long long resets;            /* patches reset to their initial u (src/patch.cpp:185-194) */
  unsigned long long start_oob;         /* patches whose start was out of bounds (src/patch.cpp:131-138) */
  unsigned long long kernel_launches;   /* kernels launched by the last run */
  unsigned long long sampled_converge;  /* of sampled_iterations: those of k_converge (the first passes of every patch) ... */
  unsigned long long sampled_walk;      /* ... and of the memoised walk; the remainder ran in k_stragglers */
} dis_stats;

DIS_API void dis_default_params(dis_params* p);
DIS_API const char* dis_last_error(void);
DIS_API const char* dis_version(void);
/* Development entry point (tests, sweeps): the scheduling knobs of DESIGN.md (DIS_OUT_SLICES, DIS_HANDOFF, ...) are
 * read from the environment once per process; this sets one and re-reads them.  value == NULL removes the override.
 * No knob changes results.  No counterpart in the reference. */
DIS_API int dis_tune(const char* name, const char* value);

/* Handle life cycle.  `device` is a CUDA ordinal. */
DIS_API int dis_create(const dis_params* p, int device, dis_engine** out);
DIS_API void dis_destroy(dis_engine* e);

/* ---- whole path, host buffers -----------------------------------------------------------
 * Replaces the body of the per-pair loop, src/main.cpp:135-198: divisibility pad (:139-155),
 * u8->f32 (:158-160), construct_pyramide x2 (:177-178, :12-50), the OpticalFlowClass constructor
 * (:184-189), the finest_scale epilogue (:192-196) and the crop (:198).
 * a, b: u8 grey frames h x w with row stride `stride` bytes.  flow_out: h x w x 2 fp32 (u,v). */
DIS_API int dis_run_pair_u8(dis_engine* e, const uint8_t* a, const uint8_t* b, int w, int h,
                            int stride, float* flow_out);

/* n_pairs independent pairs (BASELINE config 3).  a[i], b[i] frames as above; flow_out holds
 * n_pairs consecutive h x w x 2 fields. */
DIS_API int dis_run_batch_u8(dis_engine* e, const uint8_t* const* a, const uint8_t* const* b,
                             int n_pairs, int w, int h, int stride, float* flow_out);

/* A frame sequence, src/main.cpp:102 loop (frame_i, frame_i+1): n_frames frames stored
 * contiguously (frame_stride bytes apart) give n_frames-1 flows.  Pyramids are built once per
 * frame and shared by the two pairs that use it. */
DIS_API int dis_run_sequence_u8(dis_engine* e, const uint8_t* frames, int n_frames, int w, int h,
                                int stride, size_t frame_stride, float* flow_out);

/* Host-buffer calls of 8 pairs or more return the flows in slices (the finest level is searched in two parts and the first
 * part's flows travel home during the second).  enabled = 0 turns that off for this engine: worth it when something else
 * hides the copies, e.g. a second engine on the same GPU (dis_pool_submit_*).  Results do not change.  No counterpart in
 * the reference. */
DIS_API int dis_set_host_slices(dis_engine* e, int enabled);

/* ---- the exact OpticalFlowClass contract ---------------------------------------------------
 * Replaces OpticalFlow::OpticalFlowClass::OpticalFlowClass (include/optical_flow.hpp:43-54,
 * src/optical_flow.cpp:19-132) with draw_grid = false.  Six HOST pyramids indexed by absolute
 * scale 0..coarsest, planes (height>>s + 2*img_padding) x (width>>s + 2*img_padding) fp32.
 * Only img_first_dx/dy and img_second are read (the reference never reads the others on this
 * path); those three may be NULL.  outflow: 2*(width>>finest)*(height>>finest). */
DIS_API int dis_run_pyramids(dis_engine* e, const float* const* img_first,
                             const float* const* img_first_dx, const float* const* img_first_dy,
                             const float* const* img_second, const float* const* img_second_dx,
                             const float* const* img_second_dy, int img_padding, float* outflow,
                             int width, int height);

/* ---- construct_pyramide on the device (src/main.cpp:12-50) ---------------------------------
 * img: fp32 host frame width x height (already divisibility padded).  Writes lv_f+1 padded host
 * planes per pyramid, as the reference's float** outputs. */
DIS_API int dis_construct_pyramid(dis_engine* e, const float* img, int width, int height, int lv_f,
                                  int img_padding, float* const* img_pyr, float* const* img_dx_pyr,
                                  float* const* img_dy_pyr);

/* ---- device-resident pipeline (what bench.py's `value` times) ------------------------------
 * upload: copies n_frames u8 frames to the device and defines the pairs (first[i], second[i])
 * as indices into them.  run: pyramids + all levels for all pairs, asynchronously on the
 * handle's stream.  download: copies cropped flows (n_pairs x h x w x 2) to the host. */
DIS_API int dis_upload_frames_u8(dis_engine* e, const uint8_t* frames, int n_frames, int w, int h,
                                 int stride, size_t frame_stride, const int* first,
                                 const int* second, int n_pairs);
DIS_API int dis_run_resident(dis_engine* e);
DIS_API int dis_download_flows(dis_engine* e, float* flow_out);
DIS_API int dis_synchronize(dis_engine* e);
/* CUDA-event timing on the handle's own stream. */
DIS_API int dis_timer_start(dis_engine* e);
DIS_API int dis_timer_stop_ms(dis_engine* e, float* ms);
/* Device time spent in the inverse-search kernels during the last dis_run_resident (events around
 * each launch), and the number of such launches. */
DIS_API int dis_last_search_ms(dis_engine* e, float* ms, int* launches);
/* Same, for the launch of one level (level = scale - finest_scale). */
DIS_API int dis_last_search_level_ms(dis_engine* e, int level, float* ms);
/* Same, for one of the three kernels of a level: phase 0 = k_converge (passes 0..23, every pass
 * sampled), 1 = k_inverse_search (the memoised walk), 2 = k_stragglers. */
DIS_API int dis_last_search_phase_ms(dis_engine* e, int level, int phase, float* ms);

/* ---- introspection for the parity harness ---------------------------------------------------
 * Valid after any run; `pair` indexes the pairs of that run, `level` = scale - finest_scale. */
DIS_API int dis_num_levels(dis_engine* e);
DIS_API int dis_get_level_info(dis_engine* e, int level, dis_level_info* out);
/* PatchGrid::get_patch_ref_pos / Patch::get_u_vector / counter_iter for every patch of a level
 * (include/patch_grid.hpp:28-30, include/patch.hpp:41-43).  Any output may be NULL. */
DIS_API int dis_get_patches(dis_engine* e, int pair, int level, float* ref_xy, float* u_xy,
                            int* iters);
/* The level's dense flow (flows[s] of src/optical_flow.cpp:86-90), height x width x 2. */
DIS_API int dis_get_level_flow(dis_engine* e, int pair, int level, float* flow);
/* Padded pyramid planes of one frame of the last run: which = 0 image, 1 dx, 2 dy. */
DIS_API int dis_get_plane(dis_engine* e, int frame, int scale, int which, float* out);
DIS_API int dis_get_stats(dis_engine* e, dis_stats* out);

/* ---- PatchGrid-level API (include/patch_grid.hpp:16-30) --------------------------------------
 * A grid is one level of one pair with host-pointer inputs, mirroring the reference's methods
 * one to one; used by the C++ mirror classes in include/dis_b200.hpp. */
typedef struct dis_grid dis_grid;
DIS_API int dis_grid_create(dis_engine* e, int width, int height, int img_padding, dis_grid** out);
/* A grid of ONE patch centred at (ref_x, ref_y): the unit behind the Patch mirror class
 * (include/patch.hpp:33-43; Patch::init_patch takes the reference point explicitly).  The
 * reference rounds the point for the template gather (src/patch.cpp:52-53) and only ever passes
 * integer points; non-integer points are rejected with DIS_ERR_PARAM. */
DIS_API int dis_grid_create_single(dis_engine* e, int width, int height, int img_padding, float ref_x,
                                   float ref_y, dis_grid** out);
DIS_API void dis_grid_destroy(dis_grid* g);
DIS_API int dis_grid_init(dis_grid* g, const float* img_first, const float* img_first_dx,
                          const float* img_first_dy);                     /* init_grid */
DIS_API int dis_grid_set_target(dis_grid* g, const float* img_second, const float* img_second_dx,
                                const float* img_second_dy);              /* set_target_image */
DIS_API int dis_grid_init_from_prev_flow(dis_grid* g, const float* flow_prev); /* patch_init_from_prev_flow */
DIS_API int dis_grid_set_patch_init(dis_grid* g, int i, float ux, float uy);   /* Patch::inverse_search(init) */
DIS_API int dis_grid_inverse_search(dis_grid* g);                         /* inverse_search */
DIS_API int dis_grid_densify(dis_grid* g, float* dense_flow);             /* densification_and_create_dance_flow */
DIS_API int dis_grid_num_patches(dis_grid* g);                            /* get_num_all_patch */
DIS_API int dis_grid_get_patch(dis_grid* g, int i, float* ref_xy, float* u_xy); /* get_patch_ref_pos / get_u_vector */

/* ---- "next" rows of SURVEY.md §8f ---------------------------------------------------------- */
/* Middlebury .flo writer/reader, src/IO_flow.cpp:10-98 (tag "PIEH", int32 w, h, interleaved f32). */
DIS_API int dis_write_flo(const char* path, const float* flow, int w, int h);
DIS_API int dis_read_flo(const char* path, float* flow, int w, int h);
DIS_API int dis_read_flo_size(const char* path, int* w, int* h);

/* ---- one large frame tiled by grid rows over several GPUs (BASELINE config 5) ------
 * Replaces the same per-pair path as dis_run_pair_u8 (src/main.cpp:135-198) when the pair is split
 * over N engines (one per GPU).  Every engine gets the whole pair; per level (coarsest first,
 * level = scale - finest_scale) it searches the grid rows [row0, row1) (PatchGrid::inverse_search,
 * src/patch_grid.cpp:99-106, restricted to those patches), the per-patch flows
 * (Patch::get_u_vector, [num_patches][2] fp32, patch id = x*nph + y) are exchanged by the caller
 * through DEVICE buffers on the engine's GPU (e.g. an NCCL all-gather), and the level is densified
 * (src/patch_grid.cpp:121-182) redundantly on every engine.  Bit-identical to dis_run_pair_u8. */
DIS_API int dis_tiled_begin(dis_engine* e, const uint8_t* a, const uint8_t* b, int w, int h, int stride);
DIS_API int dis_tiled_search(dis_engine* e, int level, int row0, int row1);
DIS_API int dis_tiled_get_u(dis_engine* e, int level, void* device_dst);
DIS_API int dis_tiled_set_u(dis_engine* e, int level, const void* device_src);
DIS_API int dis_tiled_densify(dis_engine* e, int level);
DIS_API int dis_tiled_finish(dis_engine* e, float* flow_out);
/* The same run with a HALO exchange instead of the all-gather, for engines of ONE process (dis_pool_run_pair_tiled_u8
 * drives these; all calls are asynchronous, cross-GPU ordering is by CUDA events, nothing synchronises between levels):
 * fetch_u_rows copies grid rows [row0, row1) of `src`'s per-patch flows of the level into e's (src on the same or on a
 * peer GPU with peer access enabled: the copying SMs read over NVLink) once src has searched them; densify_rows
 * densifies pixel rows [y0, y1) of the level only (src/patch_grid.cpp:121-182 for those rows — the caller makes sure
 * the covering patches, ceil(ps / (2 steps)) + 1 grid rows beyond the band, are present); finish_rows crops and
 * downloads rows [y0, y1) of the result straight into their place in flow_out; dis_tiled_end waits and closes the run. */
DIS_API int dis_tiled_fetch_u_rows(dis_engine* e, int level, dis_engine* src, int row0, int row1);
DIS_API int dis_tiled_densify_rows(dis_engine* e, int level, int y0, int y1);
DIS_API int dis_tiled_finish_rows(dis_engine* e, float* flow_out, int y0, int y1);
DIS_API int dis_tiled_end(dis_engine* e);
/* device time (CUDA events on e's stream) of the last halo-exchange run on this engine: pyramids + all levels, from
 * after the upload was enqueued to before the download */
DIS_API int dis_tiled_last_ms(dis_engine* e, float* ms);

/* ---- several GPUs of one box behind one handle ------------------------------------------------------------------
 * The reference's unit of independent work is one iteration of its per-pair loop (src/main.cpp:102; the body shares no
 * state).  A pool holds one engine and one host thread per listed CUDA device; a batch / sequence is cut into
 * CONTIGUOUS blocks of pairs, one per device (sizes differ by at most one, earlier devices get the longer blocks;
 * neighbouring blocks of a sequence share one frame so that every device builds each pyramid once), and every block runs
 * through dis_run_batch_u8 / dis_run_sequence_u8 on its device concurrently.  No collective and no peer traffic: each
 * GPU writes its flows straight into the caller's buffer.  Results are bit-identical to one engine.  A device may be
 * listed more than once (two engines on one GPU).  Worker threads bind to the CPUs of their GPU's NUMA node
 * (DIS_POOL_NO_AFFINITY=1 disables); dis_pool_host_alloc allocates pinned host memory from such a thread. */
typedef struct dis_pool dis_pool;
DIS_API int dis_pool_create(const dis_params* p, const int* devices, int n_devices, dis_pool** out);  /* devices NULL: 0..n-1 */
DIS_API void dis_pool_destroy(dis_pool* pool);
DIS_API int dis_pool_size(const dis_pool* pool);
DIS_API int dis_pool_device(const dis_pool* pool, int worker, int* device, int* numa_cpus);  /* numa_cpus: CPUs the thread is bound to (0: not bound) */
DIS_API const char* dis_pool_last_error(void);
/* block [lo, hi) of part `part` of `parts`: the sharding rule itself (pure; also used by the host-side tests) */
DIS_API void dis_pool_shard(int n_pairs, int parts, int part, int* lo, int* hi);
DIS_API int dis_pool_run_batch_u8(dis_pool* pool, const uint8_t* const* a, const uint8_t* const* b, int n_pairs, int w,
                                  int h, int stride, float* flow_out);                       /* BASELINE config 3 */
DIS_API int dis_pool_run_sequence_u8(dis_pool* pool, const uint8_t* frames, int n_frames, int w, int h, int stride,
                                     size_t frame_stride, float* flow_out);                  /* BASELINE config 2 over N GPUs */
/* Asynchronous form of the two calls above (the reference's loop, src/main.cpp:102, has no counterpart: it is sequential).
 * A submission is ONE whole dis_run_sequence_u8 / dis_run_batch_u8 on ONE engine of the pool — the idle one, else the one
 * with the shortest queue — and returns a ticket at once; the caller's buffers belong to the pool until dis_pool_wait
 * returns for that ticket (its status is the call's; stats / worker may be NULL).  List a GPU TWICE in dis_pool_create and
 * keep two submissions in flight: the uploads and downloads of one run under the kernels of the other (bench.py's `e2e`).
 * Results are bit-identical to the blocking calls.  Tickets are waited for in any order, each once; dis_pool_wait_all
 * drains every outstanding ticket and returns the first failure.  One host thread submits and waits. */
DIS_API int dis_pool_submit_sequence_u8(dis_pool* pool, const uint8_t* frames, int n_frames, int w, int h, int stride,
                                        size_t frame_stride, float* flow_out, int* ticket);
DIS_API int dis_pool_submit_batch_u8(dis_pool* pool, const uint8_t* const* a, const uint8_t* const* b, int n_pairs, int w,
                                     int h, int stride, float* flow_out, int* ticket);
DIS_API int dis_pool_wait(dis_pool* pool, int ticket, dis_stats* stats, int* worker);
DIS_API int dis_pool_wait_all(dis_pool* pool);
/* ONE large pair split over the pool's GPUs by bands of grid rows with a per-level halo exchange over NVLink peer
 * memory (BASELINE config 5).  Every GPU gets both frames and builds the pyramids redundantly (< 1 % of
 * the work); per level it searches only the patches of its band, pulls the ceil(ps / (2 steps)) + 1 boundary grid rows
 * it needs from its neighbours, densifies only the pixel rows its next-finer band is initialised from (finest level:
 * its band of the output) and downloads only that band.  Bit-identical to dis_run_pair_u8.  finest_scale must be 0. */
DIS_API int dis_pool_run_pair_tiled_u8(dis_pool* pool, const uint8_t* a, const uint8_t* b, int w, int h, int stride,
                                       float* flow_out);
/* device time of the last dis_pool_run_pair_tiled_u8: the slowest GPU's pyramids + levels (CUDA events), i.e. the run
 * without its host<->device copies */
DIS_API int dis_pool_last_tiled_device_ms(dis_pool* pool, double* max_ms);
DIS_API int dis_pool_get_stats(dis_pool* pool, dis_stats* sum);                              /* summed over the devices of the last call */
DIS_API int dis_pool_last_ms(dis_pool* pool, double* wall_ms, double* per_worker_ms);        /* host wall clock of the last call */
DIS_API int dis_pool_host_alloc(dis_pool* pool, int worker, size_t bytes, void** out);       /* pinned, NUMA-local to that worker's GPU */
DIS_API int dis_pool_host_free(dis_pool* pool, void* ptr);

/* Middlebury colour coding, draw_optical_flow(flow, dst, maxmotion) of src/color_coding.cpp:83-117
 * (called at src/main.cpp:200-201 with maxmotion = -1: the range is the field's own maximum).
 * flow: h x w x 2 fp32 host; bgr_out: h x w x 3 u8 host (B,G,R as cv::Mat CV_8UC3). */
DIS_API int dis_color_code(dis_engine* e, const float* flow, int w, int h, float maxmotion, uint8_t* bgr_out);
/* Same for the (device-resident) cropped flow of pair `pair` of the last run.
 * NOTE "last run": dis_run_batch_u8 / dis_run_sequence_u8 work in chunks of at most 128 pairs; after a longer call
 * the device holds only the LAST chunk, and `pair` (here, in dis_get_patches and in dis_get_level_flow) indexes that
 * chunk: pair 0 is pair 128 * floor((n_pairs - 1) / 128) of the call.  Colour a long sequence from the host flows
 * with dis_color_code (tools/dis_sequence.cpp does). */
DIS_API int dis_color_code_pair(dis_engine* e, int pair, float maxmotion, uint8_t* bgr_out);

/* ---- colour coding without any allocation ------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arithmetic and
 * byte-identical output as dis_color_code, but d_flow (h x w x 2 fp32) and d_bgr_out (h x w x 3 u8) are the CALLER's
 * DEVICE buffers on CUDA device `device`, the work is enqueued on `cuda_stream` (a cudaStream_t; NULL = the default
 * stream) and the call returns without synchronising.  It needs no engine handle and calls no cudaMalloc / cudaFree /
 * allocating routine: the scratch of the max-|u| reduction is a fixed 33 KB __device__ array inside the library (one
 * per GPU that loads it, held until exit).  Limit: w * h <= DIS_NOALLOC_MAX_PIXELS (33 554 432, e.g. 7680 x 4320);
 * larger fields are rejected with DIS_ERR_PARAM (use dis_color_code).  Every call on a GPU shares that one array:
 * calls that overlap on the GPU (different streams or threads) race; calls ordered on one stream do not. */
#define DIS_NOALLOC_PIXELS_PER_PARTIAL 4096
#define DIS_NOALLOC_MAX_PIXELS (8192 * DIS_NOALLOC_PIXELS_PER_PARTIAL)
DIS_API int dis_color_code_device_noalloc(int device, const float* d_flow, int w, int h, float maxmotion,
                                          uint8_t* d_bgr_out, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* DIS_B200_H */
