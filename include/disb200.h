/* disb200 — C-ABI of the B200-native DIS optical-flow / depth-from-stereo hot path.
 *
 * Drop-in boundary for lordnn/OF_DIS (reference paths are relative to the upstream tree):
 *   - the parameter block mirrors the 20 positional CLI parameters of run_dense.cpp:272-292
 *     plus the two compile-time switches SELECTMODE / SELECTCHANNEL (CMakeLists.txt:38-61);
 *   - disb200_ofclass      replaces the OFC::OFClass constructor (oflow.h:99-126; only call
 *                          site run_dense.cpp:392-401): same padded host pyramids in, same
 *                          low-resolution interleaved flow out;
 *   - disb200_run_batch*   replace main() of run_dense.cpp:297-415 between imread and the file
 *                          writers: dimension padding, ConstructImgPyramide, OFClass, final
 *                          upsample and crop, for a batch of independent frame pairs;
 *   - disb200_save_flo / disb200_save_pfm replace SaveFlowFile / SavePFMFile
 *                          (run_dense.cpp:17-82).
 * All functions return 0 on success and a non-zero DISB200_E* code otherwise (the reference
 * has no error path: void ctor, exit(1) on OOM — image.cpp:25-41); disb200_last_error()
 * returns a thread-local message.  No function ever falls back to a CPU implementation:
 * without a CUDA device every compute entry point fails with DISB200_ECUDA.
 */
#ifndef DISB200_H_
#define DISB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  DISB200_OK = 0,
  DISB200_EINVAL = 1,   /* bad dimensions / parameters */
  DISB200_ECUDA = 2,    /* CUDA runtime error or no device */
  DISB200_ENOMEM = 3,   /* device or host allocation failed */
  DISB200_EUNSUPPORTED = 4, /* parameter combination the kernels do not cover */
  DISB200_EIO = 5
};

enum { DISB200_MODE_OF = 1, DISB200_MODE_DE = 2 };       /* SELECTMODE    (patch.h:28-34) */
enum { DISB200_SOR_LEXICOGRAPHIC = 0, DISB200_SOR_REDBLACK = 1 };

/* run_dense.cpp:272-292 in CLI order, then the compile-time variant selectors. */
typedef struct DisParams {
  int sc_f;          /* coarsest scale                                   (argv 4) */
  int sc_l;          /* finest scale                                     (argv 5) */
  int max_iter;      /* inverse-search iterations, upper bound            (6) */
  int min_iter;      /*                            lower bound            (7) */
  float dp_thresh;   /* min rate of change of the step norm               (8) */
  float dr_thresh;   /* min rate of change of the residual                (9) */
  float res_thresh;  /* stop when mean |residual| falls below            (10) */
  int patchsz;       /* patch edge length in px; also the level padding   (11) */
  float patove;      /* patch overlap in [0,1)                            (12) */
  int usefbcon;      /* forward-backward merging                          (13) */
  int patnorm;       /* mean-normalise patches                            (14) */
  int costfct;       /* 0 L2, 1 L1, 2 pseudo-Huber, 10 NCC                (15) */
  int usetvref;      /* variational refinement on/off                     (16) */
  float tv_alpha;    /* smoothness weight                                 (17) */
  float tv_gamma;    /* gradient-constancy weight                         (18) */
  float tv_delta;    /* colour-constancy weight                           (19) */
  int tv_innerit;    /* fixed-point iterations (x (level+1))              (20) */
  int tv_solverit;   /* SOR sweeps per fixed-point iteration              (21) */
  float tv_sor;      /* SOR omega                                         (22) */
  int verbosity;     /* 0 silent, 1 total time, 2 per-scale times         (23) */
  int mode;          /* DISB200_MODE_OF | DISB200_MODE_DE  (SELECTMODE)        */
  int channels;      /* 1 grey | 3 BGR interleaved        (SELECTCHANNEL)     */
  int sor_mode;      /* DISB200_SOR_LEXICOGRAPHIC (parity with solver.c:83-433) | _REDBLACK */
  int gradmag;       /* 1: SELECTCHANNEL == 2 — the grey frame is replaced by the magnitude of its Sobel/8
                        gradient before the pyramid is built (run_dense.cpp:139-148); needs channels == 1.
                        0: the frame is used as given (SELECTCHANNEL 1 | 3) */
} DisParams;

typedef struct DisHandle DisHandle;

/* Schedule selection (every choice computes the same result; the automatic one is what the benchmarks run).  Read when a
 * handle is created (and by the one-shot stage entry points); 0 in a field = automatic.  Used by the tests to force each
 * SOR schedule against the oracle and by the A/B measurements in DESIGN.md. */
typedef struct DisTuning {
  int streams;       /* sub-batches of one run_batch call on their own streams (auto: 2 from 1184 pairs up) */
  int host_chunk;    /* pairs per pipelined chunk of the host-pointer entry points (auto: n / 8, at least 16) */
  int tv_schedule;   /* refinement of a level: 1 fused (one CTA per pair runs the whole level), 2 split with system_kernel and
                        the stand-alone one-warp / banded SOR, 3 split with the general CTA wavefront SOR, 4 split with the
                        system assembled inside the one-warp SOR kernel (an experiment that measured 2.9x slower; never automatic),
                        5 split with the team SOR (one warp per 32-row band, one CTA per pair: what 0 chooses for levels of 33 - 96
                        rows; 2 keeps one warp per pair there) */
  int sorb_max;      /* largest n * bands the banded SOR takes (auto and upper bound: the number of co-resident CTAs) */
  int patch_map;     /* 1: value-interleaved lane mapping of the inverse search where the block mapping is the default,
                        and direct global tap loads where the 8x8 grey search stages its tap windows in shared memory */
  int antiphase;     /* 1: with two sub-batches, run their refinements in anti-phase (one assembles while the other solves) */
} DisTuning;
int disb200_set_tuning(const DisTuning* t);   /* NULL restores the automatic choices */

/* Operating points 1..4 of run_dense.cpp:226-268 (anything else selects 2), including
 * AutoFirstScaleSelect (run_dense.cpp:181-184) on the unpadded image width. */
int disb200_operating_point(int op, int width, int mode, int channels, DisParams* out);

/* Dimension padding of run_dense.cpp:299-312: W,H = dimensions padded to a multiple of
 * 2^sc_f, padl/padt = columns/rows added on the left/top (floor(pad/2)). */
int disb200_padded_dims(int w0, int h0, int sc_f, int* W, int* H, int* padl, int* padt);

/* Number of devices visible to the library (0 when there is no GPU or no driver). */
int disb200_device_count(void);

/* One handle per (GPU, stream set): owns every device workspace for batches of up to
 * max_batch pairs of w0 x h0 x channels u8 frames.  No allocation happens in run_batch*. */
int disb200_create(const DisParams* params, int w0, int h0, int max_batch, int device, DisHandle** out);
int disb200_destroy(DisHandle* h);

/* Host-side batch: imgA[i], imgB[i] are h0*w0*channels u8 (row-major, channels interleaved,
 * BGR like cv::imread), out[i] receives h0*w0*nop f32 (nop = 2 for OF: u,v interleaved; 1 for
 * DE).  Includes the host->device and device->host copies; returns when out[] is complete.
 * n may exceed max_batch (processed in chunks). */
int disb200_run_batch(DisHandle* h, const uint8_t* const* imgA, const uint8_t* const* imgB, int n, float* const* out);

/* Same, but frames and results are contiguous in pinned or pageable host memory:
 * A, B: n*h0*w0*channels u8; out: n*h0*w0*nop f32.  Chunks are double-buffered so copies
 * overlap compute. */
int disb200_run_batch_host(DisHandle* h, const uint8_t* A, const uint8_t* B, int n, float* out);

/* The OFClass cut with u8 frames in: same pipeline, but the result is what OFC::OFClass writes to
 * `outflow` (oflow.h:105, run_dense.cpp:385-401) — the flow of level sc_l, (H >> sc_l) x (W >> sc_l) x nop
 * f32 per pair in level-sc_l pixel units — i.e. without the `flow *= 2^sc_l; cv::resize; crop` of
 * run_dense.cpp:407-415.  For consumers that keep working on the GPU grid or upsample on their side: at
 * op2 it moves 64 x fewer bytes to the host than the full-resolution flow, which is what bounds
 * disb200_run_batch_host (PCIe).  disb200_lowres_dims reports the per-pair shape. */
int disb200_run_batch_host_lowres(DisHandle* h, const uint8_t* A, const uint8_t* B, int n, float* out_low);
int disb200_lowres_dims(const DisHandle* h, int* w_l, int* h_l, int* nop);

/* Device-resident batch (n <= max_batch): dA, dB, dOut are device pointers laid out as above.
 * Asynchronous on the handle's stream; call disb200_sync before reading dOut. */
int disb200_run_batch_device(DisHandle* h, const uint8_t* dA, const uint8_t* dB, int n, float* dOut);
int disb200_sync(DisHandle* h);

/* Video: nframes consecutive frames (contiguous, each h0*w0*channels u8) -> nframes-1 flows, pair i =
 * frames (i, i+1).  The front end SURVEY.md 8(f) ranks first: every frame's pyramid is built once and serves
 * both pairs it belongs to (the reference would rebuild it, run_dense.cpp:344-345).  Results are identical to
 * run_batch* on the explicit pairs.  _device: nframes-1 <= max_batch, asynchronous like run_batch_device. */
int disb200_run_sequence_device(DisHandle* h, const uint8_t* dFrames, int nframes, float* dOut);
int disb200_run_sequence_host(DisHandle* h, const uint8_t* frames, int nframes, float* out);
int disb200_run_sequence_host_lowres(DisHandle* h, const uint8_t* frames, int nframes, float* out_low);

/* Video with the temporal warm start of oflow.cpp:213-217 (the `initflow` argument; its only caller ships commented out,
 * run_dense.cpp:354-379): pair i starts its coarsest level from pair i-1's result instead of zero - the level-sc_l flow of
 * pair i-1, box-averaged (INTER_AREA, run_dense.cpp:372) down to the (H >> (sc_f+1)) x (W >> (sc_f+1)) grid OFClass expects
 * and rescaled to that grid's units; pair 0 starts cold.  Pyramids are shared like in run_sequence_*; the pairs run one
 * after another (each depends on the one before), so this is the latency path.  Results differ from run_sequence_* (a
 * different starting point); they equal OFClass called pair by pair with that initflow. */
int disb200_run_sequence_warm_device(DisHandle* h, const uint8_t* dFrames, int nframes, float* dOut);
int disb200_run_sequence_warm_host(DisHandle* h, const uint8_t* frames, int nframes, float* out);

/* CUDA-event stopwatch on the handle's stream (the stream every kernel is launched on). */
int disb200_timer_begin(DisHandle* h);
int disb200_timer_end(DisHandle* h, float* elapsed_ms);
/* Kernels launched by the most recent run_batch* call on this handle. */
int disb200_last_launch_count(const DisHandle* h);
/* Device time in ms the most recent run spent in stage s (0 pyramid, 1 inverse search,
 * 2 densification, 3 variational refinement as a whole, 4 upsample, and inside 3: 5 the SOR kernel
 * (or the fused per-level kernel of small batches), 6 system_kernel, 7 smooth_kernel2); only recorded while
 * disb200_set_stage_timing(h, 1) is on (it serialises the stream with events). */
int disb200_set_stage_timing(DisHandle* h, int on);
int disb200_stage_ms(const DisHandle* h, int stage, float* ms);

/* The per-scale split the reference prints at verbosity 2 (oflow.cpp:292-298): device time of the inverse search (the
 * reference's pconst + pinit + poptim: one kernel here), the densification (cflow) and the refinement (tvopt) of pyramid
 * level `level` in the most recent run with stage timing on, and the level's patch count (#p). */
int disb200_scale_ms(const DisHandle* h, int level, float* search_ms, float* densify_ms, float* refine_ms, int* npatches);

/* Intermediate results of the most recent run (parity tests): copies to host
 *   kind 0/1/2 : padded pyramid image / dx / dy of frame `which` (0 = A, 1 = B) at `level`
 *   kind 3     : dense flow of direction `which` (0 fw, 1 bw) at `level` after refinement
 *   kind 4/5/6 : per-patch p_iter / pt_iter / pweight of direction `which` at level sc_l
 * nbytes must equal the size of the object; see INTEGRATION.md for shapes. */
int disb200_fetch(DisHandle* h, int kind, int which, int level, int pair, void* dst, size_t nbytes);

/* OFClass-level entry (oflow.h:99-126), one pair, padded host pyramids valid for levels
 * sc_l..sc_f; outflow: (H>>sc_l)*(W>>sc_l)*nop f32; initflow optional (may be NULL). */
int disb200_ofclass(const DisParams* params, const float* const* im_ao, const float* const* im_ao_dx,
                    const float* const* im_ao_dy, const float* const* im_bo, const float* const* im_bo_dx,
                    const float* const* im_bo_dy, int W, int H, float* outflow, const float* initflow);

/* The same on an existing handle (no allocation per call: what a maintainer who replaces the OFClass constructor inside a
 * loop over frames wants).  W, H are the handle's padded dimensions (disb200_padded_dims of its w0, h0);
 * uses pair slot 0 of the handle's workspace and returns when outflow is complete. */
int disb200_ofclass_handle(DisHandle* h, const float* const* im_ao, const float* const* im_ao_dx, const float* const* im_ao_dy,
                           const float* const* im_bo, const float* const* im_bo_dx, const float* const* im_bo_dy,
                           float* outflow, const float* initflow);

/* VarRefClass-level entry (refine_variational.h:35-37): refine `flow` (h_l*w_l*nop, in place)
 * on one level given the padded level images of both frames. */
int disb200_varref_level(const DisParams* params, const float* im_a, const float* im_b, int w_l, int h_l,
                         int level, int camlr, float* flow);

/* PatGridClass::AggregateFlowDense-level entry (patchgrid.cpp:213-397, called at oflow.cpp:259-267): densify the patch
 * results of one w_l x h_l level.  p_iter [npatch][nop], pweight [npatch][novals] of the grid itself; cg_* = p_iter,
 * pt_iter [npatch][2] (end positions) and pweight of the complementary grid for forward-backward merging, or all NULL.
 * flow: h_l*w_l*nop f32 out.  npatch as PatGridClass computes it (patchgrid.cpp:43-46); returned through *npatch_out
 * when not NULL so that callers can size the arrays (call with all data pointers NULL to query). */
int disb200_aggregate_dense(const DisParams* params, int w_l, int h_l, const float* p_iter, const float* pweight,
                            const float* cg_p_iter, const float* cg_pt_iter, const float* cg_pweight, float* flow,
                            int* npatch_out);

/* Allocation-free variant of disb200_aggregate_dense.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Calls no cudaMalloc / cudaFree / allocating routine.  All array arguments are DEVICE pointers on the current
 * device, laid out as for disb200_aggregate_dense (d_cg_* all NULL: no forward-backward merge); the work is enqueued on
 * `stream` (a cudaStream_t; NULL = the default stream) and the call returns without synchronising.  The binning scratch of
 * the merge lives in a static device array of DISB200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until
 * exit: a level needing more - 3 * cells + npatch ints, cells = ((w_l+2psz)/psz+2) * ((h_l+2psz)/psz+2) - is refused
 * with DISB200_EINVAL (1 MiB covers e.g. 1920 x 1080 with 8 x 8 patches at 0.4 overlap: 231 k of the 262 k ints).  Every call shares
 * the array: calls that overlap on one GPU (different streams, or threads) race on it - serialise them. */
#define DISB200_STATIC_SCRATCH_BYTES (1 << 20)
int disb200_aggregate_dense_static(const DisParams* params, int w_l, int h_l, const float* d_p_iter, const float* d_pweight,
                                   const float* d_cg_p_iter, const float* d_cg_pt_iter, const float* d_cg_pweight,
                                   float* d_flow, void* stream);

/* ---- One frame pair tiled spatially over several GPUs of one node (4K / 8K frames) ----
 * One host thread, ndev devices with peer access (NVLink).  Small pyramid levels are computed redundantly on every
 * device; levels taller than 96 rows are cut into strips of whole 32-row bands: patches are owned by the row of their
 * reference point and all-gathered after the inverse search, densification and refinement run per strip with locally
 * recomputed halo rows, the lexicographic SOR runs as one wavefront across all devices (boundary rows stored straight
 * into the neighbour's memory), the refined flow rows are all-gathered for the next level, and every device upsamples and
 * returns its rows of the result.  Bit-identical to disb200_run_batch* on one device.  `devices` may repeat a device id
 * (several strips on one GPU; what the single-GPU tests use).  Same parameter domain as disb200_create. */
typedef struct DisTiled DisTiled;
int disb200_tiled_create(const DisParams* params, int w0, int h0, int ndev, const int* devices, DisTiled** out);
int disb200_tiled_destroy(DisTiled* t);
/* A, B: h0*w0*channels u8 on the host, out: h0*w0*nop f32 on the host; returns when out is complete. */
int disb200_tiled_run_host(DisTiled* t, const uint8_t* A, const uint8_t* B, float* out);
/* Latency of the device part: reruns the pipeline `reps` times on the frames the last run_host call left on the devices
 * (results stay there) and reports the mean device time per pair, measured between two all-device barriers. */
int disb200_tiled_time(DisTiled* t, int reps, float* ms_per_pair);
/* How `level` is split: *tiled = 0 (replicated on every device) | 1; rows / patch_rows: ndev + 1 boundaries each. */
int disb200_tiled_plan(const DisTiled* t, int level, int* tiled, int* rows, int* patch_rows);
int disb200_tiled_last_launch_count(const DisTiled* t);

/* File writers with the reference's framing (run_dense.cpp:17-82). */
int disb200_save_flo(const char* path, const float* flow, int w, int h, int nc);
int disb200_save_pfm(const char* path, const float* disp, int w, int h);
/* ReadFlowFile (run_dense.cpp:84-128): the caller states nc (1 depth, 2 flow, 4 scene flow; the file does not store
 * it); *data is malloc'ed (release with disb200_free).  Short, long or mistagged files give DISB200_EIO. */
int disb200_read_flo(const char* path, int nc, float** data, int* w, int* h);
/* Binary PGM (P5) / PPM (P6) reader standing in for cv::imread (run_dense.cpp:209-210):
 * want_channels 1 -> grey, 3 -> BGR interleaved.  *data is malloc'ed; release with disb200_free. */
int disb200_read_pnm(const char* path, int want_channels, uint8_t** data, int* w, int* h);
/* The same for PGM / PPM / PNG, chosen by the file's magic (PNG: non-interlaced, any colour type and bit
 * depth, inflated with zlib; colour -> grey as OpenCV's PNG decoder does it for IMREAD_GRAYSCALE). */
int disb200_read_image(const char* path, int want_channels, uint8_t** data, int* w, int* h);
void disb200_free(void* p);

/* Asynchronous writers: SaveFlowFile / SavePFMFile (run_dense.cpp:17-82) on a pool of host threads, so
 * that the thread driving the GPU never waits for the disk.  submit copies the array (it may be reused at
 * once) and blocks only while more than max_queued_bytes (0 = 1 GiB) wait to be written; wait returns the
 * first error of the files submitted since the last wait. */
typedef struct DisWriter DisWriter;
int disb200_writer_create(int threads, size_t max_queued_bytes, DisWriter** out);
int disb200_writer_submit_flo(DisWriter* w, const char* path, const float* flow, int width, int height, int nc);
int disb200_writer_submit_pfm(DisWriter* w, const char* path, const float* disp, int width, int height);
int disb200_writer_wait(DisWriter* w);
void disb200_writer_destroy(DisWriter* w);

const char* disb200_last_error(void);
const char* disb200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* DISB200_H_ */
