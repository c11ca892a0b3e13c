/* mcpt.h — C ABI of the CUDA hot path (libmcpt_cuda.so, sm_100a).
 *
 * This library stands where the reference's OpenCL layer stands: OpenCLApp.{h,cpp} (frame
 * orchestrator), oclbasic.{h,cpp} (device runtime) and kernels/*.cl.  Every entry point names the
 * reference interface it replaces.  Records crossing the boundary are the PODs of mcpt_types.h,
 * byte-identical to objdef.h:21-99, so a reference-side caller passes its std::vector storage
 * unchanged.  No exceptions cross the ABI: 0 = OK, negative = error, text via mcpt_last_error().
 * A context is thread-compatible (one caller at a time), there are no globals, and there is NO
 * CPU fallback: every call fails with MCPT_ERR_CUDA when no sm_100 device is usable.
 *
 *   reference                                              replacement
 *   OpenCL::init()              OpenCLApp.cpp:36-55         mcpt_create + mcpt_upload_scene + mcpt_set_camera
 *   SceneBuild::buildScene()    scenebuild.cpp:50-101,149   mcpt_upload_scene
 *   RayGeneration::init()       raygeneration.cpp:41-67     mcpt_set_camera
 *   RayGeneration::generateRay  raygeneration.cpp:29-35     mcpt_generate_rays      (kernel generateRay, rayGenerator.cl:1-31)
 *   SceneBase::intersect        scenebuild.cpp:103-133      mcpt_intersect          (kernel intersectRays, intersect.cl:1-28)
 *   SceneBase::shade            scenebuild.cpp:136-145      mcpt_shade              (kernel shade, shade.cl:75-206)
 *   ColorOut::outputColorCL     colorout.cpp:40-73          mcpt_accumulate + mcpt_display (history.cl:3-28, testkernel.cl:1-15)
 *   OpenCL::update()            OpenCLApp.cpp:57-82         mcpt_frame (stage kernels, one sample) / mcpt_render (fused wavefront)
 *   OpenCLBasic::readBuffer     oclbasic.cpp:220-222        mcpt_read_*
 *   OpenCLBasic::writeBuffer    oclbasic.cpp:224-226        mcpt_write_*
 *   outputPicture's readback    thirdpartywrapper.cpp:14-23 mcpt_read_history (then mcpt_host_write_hdr)
 */
#ifndef MCPT_H
#define MCPT_H

#include <stddef.h>
#include "mcpt_types.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mcpt_ctx mcpt_ctx;

enum {
	MCPT_OK = 0,
	MCPT_ERR_ARG = -1,      /* bad argument / call order (the reference asserts or throws a C string) */
	MCPT_ERR_CUDA = -2,     /* CUDA runtime error, no device, or wrong architecture */
	MCPT_ERR_STATE = -3     /* scene or camera not uploaded yet */
};

/* Traversal algorithms.  Both return the reference's closest hit (SURVEY.md §8c definition). */
enum {
	MCPT_TRAVERSAL_REFERENCE = 0,  /* objdef.h:240-275 as written: 64 B nodes, one box per visit, left-first DFS, no
	                                   pruning; also reproduces the raw Hit.triangleID field */
	MCPT_TRAVERSAL_FAST = 1,       /* the product: near-child-first traversal with pruning on a binary SAH tree with
	                                   two-box nodes, conservative inner tests, certified leaf-box + triangle decisions
	                                   (exact reference arithmetic whenever a decision is close), and the reference's
	                                   accept rule replayed over the near-tie candidates in canonical order (proof
	                                   sketch in csrc/cuda/mcpt_traverse.cuh) */
	MCPT_TRAVERSAL_CANONICAL = 2,  /* the reference's left-first leaf order on the two-box layout with margin-guarded
	                                   pruning; fallback of FAST and a cross-check */
	MCPT_TRAVERSAL_WIDE8 = 3       /* experimental: FAST's algorithm on 8-wide nodes with 8-bit quantised child boxes
	                                   (80 B per 8 children); same answers, measured slower than FAST on B200 (r1) */
};

typedef struct mcpt_params {
	int width, height;        /* config "width"/"height" == camera.resolution */
	int max_depth;            /* config "maxdepth"  (-D MAX_DEPTH,   scenebuild.cpp:163) */
	int max_attempt;          /* config "attempt"   (-D MAX_ATTEMPT, colorout.cpp:35)    */
	int device;               /* CUDA device ordinal (reference: devices[0], oclbasic.cpp:102) */
	int seed_mode;            /* 0: one LCG stream per pixel, seeded once (shade.cl:101, reference behaviour)
	                             1: stream re-seeded for every (pixel, sample index): order independent, used
	                                for sample-index sharding across GPUs */
	uint32_t run_seed;        /* replaces srand(time(NULL)) (scenebuild.cpp:116-119) */
	int samples_in_flight;    /* seed_mode 1: samples per wave of mcpt_render (0 = automatic); mode 0 uses 1 */
	float tmin;               /* host EPSILON 0.001f (oclbasic.h:193, scenebuild.cpp:125); 0 selects it */
	int traversal;            /* MCPT_TRAVERSAL_* used by mcpt_render / mcpt_trace (stage calls take their own) */
	int sampling;             /* MCPT_SAMPLING_REFERENCE (0): the reference's estimator, bit-exact.  MCPT_SAMPLING_NEE_MIS (1): the
	                             same image estimated with next-event estimation + MIS (mcpt_mis.cuh); needs seed_mode 1;
	                             results come from mcpt_read_sums (sum / count), history stays untouched */
	int reserved[5];
} mcpt_params;
/* MCPT_SAMPLING_NEE_MIS_RATIO (2): as NEE_MIS, and plane 6 of the sums holds an NEE + MIS estimate of the expected number
 * of counted samples instead of the raw count (whose binomial noise otherwise dominates the image), mcpt_mis.cuh */
enum { MCPT_SAMPLING_REFERENCE = 0, MCPT_SAMPLING_NEE_MIS = 1, MCPT_SAMPLING_NEE_MIS_RATIO = 2 };

typedef struct mcpt_stats {
	uint64_t rays_traced;     /* live rays that entered an intersect stage since create/reset (Mrays/s numerator) */
	uint64_t paths;           /* pixel samples finished by mcpt_render / mcpt_frame (Mpaths/s numerator) */
	uint64_t kernel_launches; /* CUDA kernels launched by this context */
	uint64_t nodes_fetched;   /* inner-node records fetched (instrumented kernels only: mcpt_trace_count, profiling) */
	uint64_t tris_fetched;    /* exact triangle tests */
	uint64_t leaves_fetched;  /* leaf records fetched (FAST; the other walks read leaf boxes from nodes) */
	uint64_t slow_rays;       /* FAST only: rays handed to the reference walk (zero direction component, far origin, ...) */
	uint64_t replayed_rays;   /* FAST only: rays re-traced by the canonical walk (candidate in the tie band, overflow) */
} mcpt_stats;

const char* mcpt_last_error(void);
int mcpt_abi_version(void);

mcpt_ctx* mcpt_create(const mcpt_params* params);          /* NULL on error */
void mcpt_destroy(mcpt_ctx* ctx);
void* mcpt_stream(mcpt_ctx* ctx);                           /* the cudaStream_t every call of ctx is ordered on */
int mcpt_synchronize(mcpt_ctx* ctx);

/* Host arrays are borrowed for the call and copied.  nodes = the canonical BVHNode[2n-1] of
 * mcpt_host_build_bvh (reference layout); the device traversal layout is derived here - for geometry the process has
 * not seen before that includes a binned-SAH tree over the same leaves, built on the context's device
 * (csrc/cuda/sah_build.cu; byte-identical to the host builder behind mcpt_derive_scene). */
int mcpt_upload_scene(mcpt_ctx* ctx, const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats,
                      const mcpt_bvh_node* nodes, int nnodes);
/* Same, with the caller's generation counter for the geometry: a non-zero tag equal to the previous upload's tag tells
 * the library that tris and nodes hold the same bytes as last time, so it re-copies the arrays (like the reference's
 * CL_MEM_COPY_HOST_PTR buffers, oclbasic.h:56-58) without comparing 80 MB to find that out.  Tag 0 = no promise. */
int mcpt_upload_scene_tagged(mcpt_ctx* ctx, const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats,
                             const mcpt_bvh_node* nodes, int nnodes, uint64_t geometry_tag);
/* Scene cache support (the reference derives nothing ahead of time - its kernels walk the uploaded
 * BVHNode[] directly, scenebuild.cpp:97-100).  mcpt_derive_scene runs the host part of mcpt_upload_scene (validation,
 * leaf records, two-box nodes, the binned-SAH tree of the fast walk) without a context or a GPU and returns it as one
 * relocatable blob (release with mcpt_free_blob) that mcpt_host_cache_store keeps next to the arrays.
 * mcpt_upload_scene_derived uploads with such a blob instead of deriving again: the blob is accepted only if its content
 * hash matches these very triangles and nodes, it was written by this library version and every link in it is in
 * range (MCPT_ERR_ARG otherwise - the caller then falls back to mcpt_upload_scene); results are identical. */
int mcpt_derive_scene(const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats,
                      const mcpt_bvh_node* nodes, int nnodes, void** blob, size_t* bytes);
/* The same blob, byte for byte, with the binned-SAH tree of the fast walk built on `device` (csrc/cuda/sah_build.cu) - what
 * mcpt_upload_scene itself does for geometry it has not seen before; *tree_ms (optional) = device time of that build. */
int mcpt_derive_scene_device(int device, const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats,
                             const mcpt_bvh_node* nodes, int nnodes, void** blob, size_t* bytes, float* tree_ms);
void mcpt_free_blob(void* blob);
int mcpt_upload_scene_derived(mcpt_ctx* ctx, const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats,
                              const mcpt_bvh_node* nodes, int nnodes, const void* blob, size_t bytes, uint64_t geometry_tag);
int mcpt_set_camera(mcpt_ctx* ctx, const mcpt_camera* cam);

/* ---- stage level: one call = one reference kernel over the W*H work items (reference buffer layouts) ---- */
int mcpt_generate_rays(mcpt_ctx* ctx);                      /* rays <- camera */
int mcpt_reset_color(mcpt_ctx* ctx);                        /* colorBuffer <- (1,1,1,1), OpenCLApp.cpp:63 */
int mcpt_intersect(mcpt_ctx* ctx, int traversal);           /* hits <- rays; dead rays untouched */
int mcpt_shade(mcpt_ctx* ctx);                              /* rays, color, seeds updated in place */
int mcpt_accumulate(mcpt_ctx* ctx);                         /* history.cl: hist/count updated, color <- running mean */
int mcpt_display(mcpt_ctx* ctx, int from_history, float* host_rgba);  /* testkernel.cl: pow(c,1/2.2), w=0 */
int mcpt_frame(mcpt_ctx* ctx);                              /* OpenCL::update(): reset, generate, maxdepth x (intersect, shade), accumulate */

/* ---- fused wavefront path (the product): spp more samples for every pixel, all on the device ---- */
int mcpt_render(mcpt_ctx* ctx, int spp);
int mcpt_render_async(mcpt_ctx* ctx, int spp);              /* enqueue only; mcpt_synchronize to wait */
int mcpt_set_first_sample(mcpt_ctx* ctx, uint32_t first);   /* seed_mode 1: global index of the next sample (seed_mode 0:
                                                               MCPT_ERR_STATE unless first == 0 - it cannot shard) */
int mcpt_get_first_sample(mcpt_ctx* ctx, uint32_t* first);  /* samples rendered so far + the first index set */
int mcpt_reset_accumulation(mcpt_ctx* ctx);                 /* hist, count, sums <- 0; seeds re-derived */

/* ---- buffer I/O (blocking, like readBuffer/writeBuffer with CL_TRUE) ---- */
int mcpt_read_rays(mcpt_ctx* ctx, mcpt_ray* dst);
int mcpt_write_rays(mcpt_ctx* ctx, const mcpt_ray* src);
int mcpt_read_hits(mcpt_ctx* ctx, mcpt_hit* dst, int32_t* closest /* may be NULL */);
int mcpt_write_hits(mcpt_ctx* ctx, const mcpt_hit* src);
int mcpt_read_color(mcpt_ctx* ctx, float* rgba);
int mcpt_write_color(mcpt_ctx* ctx, const float* rgba);
int mcpt_read_seeds(mcpt_ctx* ctx, uint32_t* dst);
int mcpt_write_seeds(mcpt_ctx* ctx, const uint32_t* src);
int mcpt_read_history(mcpt_ctx* ctx, float* hist_rgba, int32_t* counts /* may be NULL */);
/* 7 planes of W*H doubles: sum r,g,b; sum of squares r,g,b; counted (non-zero, capped) samples */
int mcpt_read_sums(mcpt_ctx* ctx, double* dst);
/* Checkpoint / resume (the reference keeps its progressive state device-resident and dumps once, colorout.cpp:23-24,
 * 49-50, 63-68): what mcpt_read_history / mcpt_read_sums / mcpt_read_seeds / mcpt_get_first_sample returned can be put
 * back into a fresh context of the same size, and rendering continues bit for bit where it stopped. */
int mcpt_write_sums(mcpt_ctx* ctx, const double* src);
int mcpt_write_history(mcpt_ctx* ctx, const float* hist_rgba, const int32_t* counts);
/* Accumulate the sums in caller-owned DEVICE memory (7*W*H doubles, e.g. the NCCL reduce buffer); NULL = own. */
int mcpt_bind_sums(mcpt_ctx* ctx, double* device_ptr);

/* Multi-GPU merge of the sample sums: dst.sums += srcs[0].sums + srcs[1].sums + ... in this order, by one
 * kernel on dst's device that reads the other contexts' accumulators in place over NVLink peer mappings (staged with
 * cudaMemcpyPeer where a peer cannot be mapped).  Stands where the reference would need one reduce per output frame;
 * synchronises every context involved.  Follow with mcpt_read_sums(dst). */
int mcpt_merge_sums(mcpt_ctx* dst, mcpt_ctx* const* srcs, int nsrc);

/* ---- intersection microbench / ad-hoc queries against the uploaded scene ---- */
/* host buffers: copies in, traces n rays, copies out (hits 48 B each; closest may be NULL) */
int mcpt_trace(mcpt_ctx* ctx, const mcpt_ray* rays, long n, mcpt_hit* hits, int32_t* closest, int traversal);
/* device-resident buffers, enqueued on mcpt_stream(ctx), no synchronisation */
int mcpt_trace_device(mcpt_ctx* ctx, const mcpt_ray* d_rays, long n, mcpt_hit* d_hits, int32_t* d_closest, int traversal);
/* instrumented build of the same kernel: counts node and triangle records fetched (device buffers) */
int mcpt_trace_count(mcpt_ctx* ctx, const mcpt_ray* d_rays, long n, int traversal);

/* Allocation-free variant of mcpt_trace.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that must not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same inputs, outputs and results as mcpt_trace(ctx, rays, n, hits, closest, MCPT_TRAVERSAL_FAST) with host
 * buffers, but the device staging buffers and the walk's work-fetch cursor live in a static device array inside the
 * library (6.25 MiB, held by every GPU that loads the library until the process exits) instead of cudaMalloc'ed scratch:
 * the call makes no cudaMalloc / cudaFree / allocating library call.  At most MCPT_TRACE_NOALLOC_MAX_RAYS rays per call
 * (MCPT_ERR_ARG beyond that, with a message naming the limit).  All calls on a device share the array, which keeps the
 * last call's contents (the parts that must start at zero are cleared by every call): calls that overlap on the GPU -
 * two host threads, or two contexts of the same device - race.  Synchronous, like mcpt_trace. */
#define MCPT_TRACE_NOALLOC_MAX_RAYS 65536
int mcpt_trace_noalloc(mcpt_ctx* ctx, const mcpt_ray* rays, long n, mcpt_hit* hits, int32_t* closest /* may be NULL */);

/* Measurement aids for bench.py (no effect on results):
 *   count_fetches: the bounce kernels of mcpt_render also count node / triangle records fetched (slower);
 *   time_bounces : CUDA events bracket every bounce-kernel launch on mcpt_stream(ctx); the summed device
 *                  time and launch count since the last call are returned by mcpt_get_bounce_time. */
int mcpt_set_profiling(mcpt_ctx* ctx, int count_fetches, int time_bounces);
int mcpt_get_bounce_time(mcpt_ctx* ctx, double* ms, uint64_t* launches);

/* Checked build (-DMCPT_CHECKED=1; tools/checked_run.sh): every node, leaf, stack, permutation, queue and path index the
 * kernels follow is compared with its bound before use; violations are counted and the first one recorded.
 * out4 = {failures, code of the first (CHK_* in mcpt_traverse.cuh), its value, its bound} for kernels run on `device`
 * so far.  The product build has no such checks and returns MCPT_ERR_STATE.  (compute-sanitizer is closed on the GPU
 * pool this was developed on; the checked build run over tools/sanitize_targets.py is what stands in for it.) */
int mcpt_check_failures(int device, uint64_t* out4);

/* NEE + MIS mode (MCPT_SAMPLING_NEE_MIS), test aid: the tabulated acceptance probability of the glossy lobe's rejection
 * loop (shade.cl:130-132) against angle(mirror direction, normal), MCPT_PHONG_ACCEPT_POINTS values uniform in [0, pi/2],
 * as the kernels interpolate it (mcpt_device.cuh).  Host only, needs no context. */
#define MCPT_PHONG_ACCEPT_POINTS 1025
int mcpt_phong_accept_table(float* out, int n);

/* Streaming-read bandwidth (GB/s, best of 5) of a scratch buffer of `bytes` read `iters` times on `device`:
 * 32 MiB measures the L2 ceiling the walk's node fetches live under, several GiB the HBM ceiling.  Needs no context. */
int mcpt_measure_read_bandwidth(int device, size_t bytes, int iters, float* gb_per_s);

int mcpt_get_stats(mcpt_ctx* ctx, mcpt_stats* out);
int mcpt_reset_stats(mcpt_ctx* ctx);
/* Diagnostic: how mcpt_render orders its bounce queues.  bit 0 = the path queue is walked in coherence-sorted order,
 * bit 1 = the shadow-ray queue of the NEE + MIS modes too; -1 = the NEE + MIS modes are still timing one wave each way
 * (the image does not depend on the order, so the library keeps whichever is faster for the scene at hand). */
int mcpt_get_queue_sorting(mcpt_ctx* ctx, int* out);

/* ---- BVH build on the device.  Replaces BVH::HLBVH<CPU>::build (BVH/hlbvh.cpp:92-200) for
 * "bvhtype": "hlbvh": host triangles in, 2*ntris-1 host nodes out, memcmp-equal to mcpt_host_build_bvh
 * (and so to the reference's array).  Needs no context.  Returns the node count or an MCPT_ERR_* code;
 * *build_ms (optional) receives the device time of the build without the two host<->device copies. */
int mcpt_build_lbvh(int device, const mcpt_triangle* tris, int ntris, mcpt_bvh_node* nodes_out, float* build_ms);

/* ---- Treelet restructuring on the device.  Replaces BVH::TreeletBVH<CPU> (BVH/treeletBVH.cpp:30-365,
 * "bvhtype": "treelet") and stands in for the reference's racy TreeletBVH<GPU> (kernels/treeletBVH.cl:230-531,
 * "bvhtype": "treeletGPU"): host nodes in (any 2N-1 array in the reference layout, e.g. from mcpt_build_lbvh),
 * host nodes out, memcmp-equal to mcpt_host_treelet.  Fails with MCPT_ERR_STATE where the reference's SAH recursion
 * would not terminate (same condition and message as the host optimiser).  nodes_out may alias nodes_in. */
int mcpt_optimize_treelet(int device, const mcpt_bvh_node* nodes_in, int nnodes, mcpt_bvh_node* nodes_out, float* build_ms);

/* ---- BVH quality metrics on the device.  Replaces BVH::TEST::SAH / EPO_GPU / LCV of the research
 * harness (bvhtest.cpp:97-110, :283-321 + kernels/EPO.cl:133-196, :324-444; config "testbvh" / "testall").
 * sah, epo and lcv hold what the reference functions return (their float results, widened); tris == NULL skips
 * EPO, camera == NULL skips LCV (the reference skips it without a "camera" entry).  epo_per_leaf / area_per_leaf
 * (optional, ntris floats, indexed by leaf slot like the kernel's buffers) receive the per-triangle values. */
typedef struct mcpt_bvh_quality {
	double sah, epo, lcv;
	double sah_exact;          /* before narrowing to float */
	double epo_area;           /* sum of the per-triangle EPO areas */
	double triangle_area;      /* sum of the triangle areas */
	double lcv_mean_leaves;    /* mean number of leaf boxes a primary ray touches */
	uint64_t lcv_rays;
	float ms_sah, ms_epo, ms_lcv;   /* device time of the three kernels */
	float pad_;
} mcpt_bvh_quality;
int mcpt_bvh_metrics(int device, const mcpt_bvh_node* nodes, int nnodes, const mcpt_triangle* tris, int ntris,
                     const mcpt_camera* camera, int width, int height, mcpt_bvh_quality* out,
                     float* epo_per_leaf, float* area_per_leaf);

#ifdef __cplusplus
}
#endif
#endif /* MCPT_H */
