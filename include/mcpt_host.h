/* mcpt_host.h — C ABI of the host data-preparation layer (libmcpt_host.so, no CUDA needed).
 *
 * These entry points stand where the reference's L0 host code stands, so a
 * maintainer can keep the reference's own callers and swap the implementation:
 *
 *   mcpt_host_load_object   <- ThirdPartyWrapper::loadObject      thirdpartywrapper.cpp:25-99
 *                              (+ normals/material id of the SceneCL ctor, scenebuild.cpp:58-62)
 *   mcpt_host_build_bvh     <- BVH::HLBVH<CPU> / TreeletBVH<CPU>  BVH/hlbvh.cpp:92-200,
 *                              selected by "bvhtype"              scenebuild.cpp:66-95
 *   mcpt_host_treelet       <- BVH::TreeletBVH<CPU>               BVH/treeletBVH.cpp:332-365
 *   mcpt_host_sah           <- BVH::TEST::SAH                     bvhtest.cpp:104-115
 *   mcpt_host_parse_camera  <- Auxiliary::parseCamera             auxiliary.cpp:20-71
 *   mcpt_host_load_config   <- Config::CONFIG                     config.cpp:70-125
 *   mcpt_host_cache_*       (new: what loadObject + the BVH build produce, kept on disk)
 *   mcpt_host_write_hdr     <- ThirdPartyWrapper::outputPicture   thirdpartywrapper.cpp:14-23
 *   mcpt_host_write_png     (new: gamma 1/2.2 8-bit view of the same buffer, testkernel.cl:9-14)
 *
 * All functions return 0 (or a count) on success and a negative value on error; the message is
 * available from mcpt_host_last_error() (thread-local).  Arrays returned through `**` parameters
 * are malloc'ed and released with mcpt_host_free().
 */
#ifndef MCPT_HOST_H
#define MCPT_HOST_H

#include <stddef.h>
#include "mcpt_types.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mcpt_host_config {
	char bvhtype[32];
	char directory[512];
	char objname[256];
	int width, height;
	int maxdepth;
	int attempt;          /* samples per pixel */
	int opencl, testbvh, testall, has_camera;
	int res_x, res_y;     /* camera.resolution */
	mcpt_camera camera;   /* valid when has_camera */
	int nobjs;            /* "testall": objname is an array of models under `directory` (config.cpp:91-95) */
	char objs[32][128];
} mcpt_host_config;

const char* mcpt_host_last_error(void);
void mcpt_host_free(void* p);

int mcpt_host_load_object(const char* directory, const char* objname, int with_normals,
                          mcpt_triangle** tris, int* ntris, mcpt_material** mats, int* nmats,
                          int** material_ids);
/* the same arrays, byte for byte, parsed by `threads` host threads (0 = all, 1 = the serial loader).
 * tinyobj (tiny_obj_loader.h:2063) and the loader above are single-threaded; 16 MB of OBJ: 0.5 s -> see README. */
int mcpt_host_load_object_mt(const char* directory, const char* objname, int with_normals, int threads,
                             mcpt_triangle** tris, int* ntris, mcpt_material** mats, int* nmats,
                             int** material_ids);
/* SURVEY §8 f-2, binary scene cache.  The reference parses the OBJ and rebuilds the BVH on every start
 * (thirdpartywrapper.cpp:25-99, scenebuild.cpp:50-101); these four calls keep that work in one file per scene:
 *   key   = content hash of the OBJ, of every MTL file it names, of `bvhtype`, with_normals and the record layouts;
 *   path  = "<directory><objname>.<bvhtype>.mcptcache" (under $MCPT_CACHE_DIR when that is set);
 *   store = Triangle[] / Material[] / material ids / BVHNode[] + an opaque blob (the traversal layout of
 *           mcpt_derive_scene, libmcpt_cuda), written to a temporary name and renamed;
 *   load  = 0 and malloc'ed copies of the same bytes on a hit; 1 on a miss (absent, other version, stale key,
 *           truncated or checksum mismatch - reason in mcpt_host_last_error()); negative on error. */
typedef struct mcpt_scene_key { uint64_t h[2]; } mcpt_scene_key;
int mcpt_host_scene_key(const char* directory, const char* objname, const char* bvhtype, int with_normals, mcpt_scene_key* key);
int mcpt_host_cache_path(const char* directory, const char* objname, const char* bvhtype, char* out, size_t cap);
int mcpt_host_cache_store(const char* path, const mcpt_scene_key* key, const mcpt_triangle* tris, int ntris,
                          const mcpt_material* mats, int nmats, const int* material_ids, const mcpt_bvh_node* nodes,
                          int nnodes, const void* derived, size_t derived_bytes);
int mcpt_host_cache_load(const char* path, const mcpt_scene_key* key, mcpt_triangle** tris, int* ntris,
                         mcpt_material** mats, int* nmats, int** material_ids, mcpt_bvh_node** nodes, int* nnodes,
                         void** derived /* may be NULL */, size_t* derived_bytes /* may be NULL */);
/* out_nodes must hold 2*n-1 nodes; returns the node count. bvhtype: "hlbvh" | "treelet" | "treeletGPU" */
int mcpt_host_build_bvh(const mcpt_triangle* tris, int n, const char* bvhtype, mcpt_bvh_node* out_nodes);
int mcpt_host_treelet(const mcpt_bvh_node* nodes_in, int nn, mcpt_bvh_node* out_nodes);
double mcpt_host_sah(const mcpt_bvh_node* nodes, int nn);
int mcpt_host_parse_camera(const char* json_text, mcpt_camera* out);
int mcpt_host_load_config(const char* path, int configid /* -1 = file's configid */, mcpt_host_config* out);
int mcpt_host_write_hdr(const char* path, int w, int h, const float* rgba);
int mcpt_host_write_png(const char* path, int w, int h, const float* rgba);

#ifdef __cplusplus
}
#endif
#endif /* MCPT_HOST_H */
