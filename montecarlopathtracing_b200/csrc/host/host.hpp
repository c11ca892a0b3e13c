// Host-side data preparation for the path tracer (C++17, no CUDA, no OpenCL).
//
// This is the "L0" layer of the reference re-written for Linux: config.json
// loading, OBJ/MTL scene loading with the reference's material classification, camera basis,
// the LBVH build whose node array is bit-identical to the reference's BVH/hlbvh.cpp, and the
// .hdr/.png writers.  Everything here feeds the C ABI in include/mcpt.h with the PODs of
// include/mcpt_types.h.
#pragma once

#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "mcpt_types.h"

namespace mcpt::host {

// ------------------------------------------------------------------ JSON (+ '#' comments)
// Minimal DOM sufficient for config.json.  The reference patches nlohmann/json's lexer to skip
// `# ... EOL` comments (json.hpp:3037-3043); this parser accepts the same superset.
struct Json {
	enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
	bool b = false;
	double num = 0.0;
	std::string str;
	std::vector<Json> arr;
	std::vector<std::pair<std::string, Json>> obj;   // insertion order kept

	bool has(const std::string& key) const;
	const Json& at(const std::string& key) const;    // throws std::runtime_error when absent
	const Json& at(size_t i) const;
	double number() const;                           // throws unless Number
	const std::string& string() const;               // throws unless String
	bool boolean() const;                            // Bool, or Number != 0
	size_t size() const { return kind == Array ? arr.size() : obj.size(); }
	static Json parse(const std::string& text);      // throws std::runtime_error with offset
};

// ------------------------------------------------------------------ config.json
// Same keys as the reference (config.cpp:70-125).  Kernel-path / platform keys
// are accepted and ignored.
struct RenderConfig {
	std::string bvhtype = "hlbvh";   // "hlbvh" | "treelet" | "treeletGPU" (treated as "treelet")
	int width = 0, height = 0;
	std::string directory, objname;
	std::vector<std::string> objs;   // "testall": objname is an array (config.cpp:91-95)
	int maxdepth = 0;
	int attempt = 0;                 // samples per pixel
	bool opencl = true;              // accepted; false makes the reference throw "Not Implemented"
	bool testbvh = false, testall = false;
	bool has_camera = false;
	Json camera;                     // position / lookat / up / fov / resolution
	int res_x = 0, res_y = 0;        // camera.resolution (ray count uses this, raygeneration.cpp:51-56)
};
RenderConfig load_config(const std::string& path, int configid_override = -1);
RenderConfig config_from_json(const Json& root, int configid_override = -1);

// ------------------------------------------------------------------ camera
// auxiliary.cpp:20-71 — identical float/double operation order.
mcpt_camera parse_camera(const Json& cam);

// ------------------------------------------------------------------ scene loading
struct Scene {
	std::vector<mcpt_triangle> triangles;   // normal + material id filled (scenebuild.cpp:58-62)
	std::vector<mcpt_material> materials;   // thirdpartywrapper.cpp:65-97 classification
	std::vector<int> material_ids;          // per triangle
	std::vector<std::string> material_names;
};
// with_normals=false leaves Triangle::normal/material_id zero exactly like loadObject() does.
Scene load_object(const std::string& directory, const std::string& objname, bool with_normals = true);
// Same result, parsed by `threads` threads (0 = all hardware threads, 1 = the serial loader above).
Scene load_object_parallel(const std::string& directory, const std::string& objname, bool with_normals, int threads);

// ------------------------------------------------------------------ binary scene cache
// One file per (OBJ + MTL content, bvhtype): the loader's and the BVH builder's outputs plus an opaque blob holding the
// traversal layout libmcpt_cuda derives from them (mcpt_derive_scene).  scene_cache.cpp.
struct SceneKey { uint64_t h[2]; };
struct CachedScene {
	std::vector<mcpt_triangle> triangles;
	std::vector<mcpt_material> materials;
	std::vector<int> material_ids;
	std::vector<mcpt_bvh_node> nodes;
	std::vector<unsigned char> derived;
	bool with_normals = true;
};
SceneKey scene_key(const std::string& directory, const std::string& objname, const std::string& bvhtype, bool with_normals);
std::string scene_cache_path(const std::string& directory, const std::string& objname, const std::string& bvhtype);
void scene_cache_store(const std::string& path, const SceneKey& key, const CachedScene& scene);      // throws on I/O errors
bool scene_cache_load(const std::string& path, const SceneKey& key, CachedScene& scene, std::string* why_not = nullptr);
// The same, into malloc'ed buffers the caller releases with free() (what the C ABI hands out): no intermediate copies.
struct RawScene {
	mcpt_triangle* triangles; mcpt_material* materials; int* material_ids; mcpt_bvh_node* nodes; unsigned char* derived;
	int ntris, nmats, nnodes; size_t derived_bytes; bool with_normals;
};
bool scene_cache_load_raw(const std::string& path, const SceneKey& key, RawScene& scene, std::string* why_not = nullptr);

// ------------------------------------------------------------------ BVH
// LBVH (BVH/hlbvh.cpp:92-200): 30-bit Morton of AABB centroids, stable sort, split at the
// highest differing bit, node numbering internal 0..N-2 / leaves N-1..2N-2.  memcmp-equal to the
// reference's node array for the same triangle list.
std::vector<mcpt_bvh_node> build_hlbvh(const std::vector<mcpt_triangle>& tris);
// Karras-2013 treelet restructuring on the CPU (BVH/treeletBVH.cpp:30-365), same slots.
std::vector<mcpt_bvh_node> optimize_treelet(const std::vector<mcpt_bvh_node>& nodes);
// SAH cost as in bvhtest.cpp:104-115 (check only).
double sah_cost(const std::vector<mcpt_bvh_node>& nodes);

// ------------------------------------------------------------------ image output
// Radiance RGBE, vertical flip, RLE scanlines as stb_image_write 1.13 emits them
// (thirdpartywrapper.cpp:14-23).  rgba = W*H float4, row 0 = image bottom.
bool write_hdr(const std::string& path, int w, int h, const float* rgba);
// 8-bit RGB PNG of pow(clamp(c),1/2.2) (testkernel.cl:9-14 display transform), same flip.
bool write_png(const std::string& path, int w, int h, const float* rgba);

}  // namespace mcpt::host
