// C exports of the host data-preparation layer (libmcpt_host.so; no CUDA dependency).
// Declared in include/mcpt_host.h.  Status convention as the device ABI: 0 = OK, negative = error,
// text via mcpt_host_last_error().
#include <cstdlib>
#include <cstring>

#include "host.hpp"
#include "mcpt_host.h"

using namespace mcpt::host;

namespace {
thread_local std::string g_err;
template <class F>
int guarded(F&& f) {
	try { return f(); }
	catch (const std::exception& e) { g_err = e.what(); return -1; }
	catch (...) { g_err = "unknown error"; return -1; }
}
template <class T>
T* dup(const std::vector<T>& v) {
	T* p = static_cast<T*>(std::malloc(v.size() * sizeof(T) + 1));
	if (!v.empty()) std::memcpy(p, v.data(), v.size() * sizeof(T));
	return p;
}
}  // namespace

extern "C" {

const char* mcpt_host_last_error(void) { return g_err.c_str(); }
void mcpt_host_free(void* p) { std::free(p); }

int mcpt_host_load_object(const char* directory, const char* objname, int with_normals,
                          mcpt_triangle** tris, int* ntris, mcpt_material** mats, int* nmats,
                          int** material_ids) {
	return guarded([&] {
		Scene s = load_object(directory, objname, with_normals != 0);
		*ntris = int(s.triangles.size());
		*nmats = int(s.materials.size());
		*tris = dup(s.triangles);
		*mats = dup(s.materials);
		*material_ids = dup(s.material_ids);
		return 0;
	});
}

int mcpt_host_load_object_mt(const char* directory, const char* objname, int with_normals, int threads,
                             mcpt_triangle** tris, int* ntris, mcpt_material** mats, int* nmats,
                             int** material_ids) {
	return guarded([&] {
		Scene s = load_object_parallel(directory, objname, with_normals != 0, threads);
		*ntris = int(s.triangles.size());
		*nmats = int(s.materials.size());
		*tris = dup(s.triangles);
		*mats = dup(s.materials);
		*material_ids = dup(s.material_ids);
		return 0;
	});
}

int mcpt_host_scene_key(const char* directory, const char* objname, const char* bvhtype, int with_normals, mcpt_scene_key* key) {
	return guarded([&] {
		SceneKey k = scene_key(directory, objname, bvhtype ? bvhtype : "hlbvh", with_normals != 0);
		key->h[0] = k.h[0]; key->h[1] = k.h[1];
		return 0;
	});
}

int mcpt_host_cache_path(const char* directory, const char* objname, const char* bvhtype, char* out, size_t cap) {
	return guarded([&] {
		std::string p = scene_cache_path(directory, objname, bvhtype ? bvhtype : "hlbvh");
		if (p.size() + 1 > cap) throw std::runtime_error("mcpt_host_cache_path: buffer too small");
		std::memcpy(out, p.c_str(), p.size() + 1);
		return 0;
	});
}

int mcpt_host_cache_store(const char* path, const mcpt_scene_key* key, const mcpt_triangle* tris, int ntris,
                          const mcpt_material* mats, int nmats, const int* material_ids, const mcpt_bvh_node* nodes,
                          int nnodes, const void* derived, size_t derived_bytes) {
	return guarded([&] {
		if (!path || !key || !tris || ntris <= 0 || (nmats > 0 && !mats) || !material_ids || (nnodes > 0 && !nodes) ||
		    (derived_bytes > 0 && !derived))
			throw std::runtime_error("mcpt_host_cache_store: null argument");
		CachedScene s;
		s.triangles.assign(tris, tris + ntris);
		s.materials.assign(mats, mats + nmats);
		s.material_ids.assign(material_ids, material_ids + ntris);
		s.nodes.assign(nodes, nodes + nnodes);
		const unsigned char* d = static_cast<const unsigned char*>(derived);
		s.derived.assign(d, d + derived_bytes);
		scene_cache_store(path, SceneKey{{key->h[0], key->h[1]}}, s);
		return 0;
	});
}

int mcpt_host_cache_load(const char* path, const mcpt_scene_key* key, mcpt_triangle** tris, int* ntris,
                         mcpt_material** mats, int* nmats, int** material_ids, mcpt_bvh_node** nodes, int* nnodes,
                         void** derived, size_t* derived_bytes) {
	return guarded([&] {
		RawScene s{};
		std::string why;
		if (!scene_cache_load_raw(path, SceneKey{{key->h[0], key->h[1]}}, s, &why)) {
			g_err = why;
			return 1;
		}
		*ntris = s.ntris; *nmats = s.nmats; *nnodes = s.nnodes;
		*tris = s.triangles; *mats = s.materials; *material_ids = s.material_ids; *nodes = s.nodes;
		if (derived) *derived = s.derived; else std::free(s.derived);
		if (derived_bytes) *derived_bytes = s.derived_bytes;
		return 0;
	});
}

int mcpt_host_build_bvh(const mcpt_triangle* tris, int n, const char* bvhtype, mcpt_bvh_node* out_nodes) {
	return guarded([&] {
		std::vector<mcpt_triangle> t(tris, tris + n);
		std::vector<mcpt_bvh_node> nodes = build_hlbvh(t);
		std::string kind = bvhtype ? bvhtype : "hlbvh";
		if (kind == "treelet" || kind == "treeletGPU") nodes = optimize_treelet(nodes);
		else if (kind != "hlbvh") throw std::runtime_error("BVH Not Implemented");   // scenebuild.cpp:76
		std::memcpy(out_nodes, nodes.data(), nodes.size() * sizeof(mcpt_bvh_node));
		return int(nodes.size());
	});
}

int mcpt_host_treelet(const mcpt_bvh_node* nodes_in, int nn, mcpt_bvh_node* out_nodes) {
	return guarded([&] {
		std::vector<mcpt_bvh_node> v(nodes_in, nodes_in + nn);
		v = optimize_treelet(v);
		std::memcpy(out_nodes, v.data(), v.size() * sizeof(mcpt_bvh_node));
		return int(v.size());
	});
}

double mcpt_host_sah(const mcpt_bvh_node* nodes, int nn) {
	std::vector<mcpt_bvh_node> v(nodes, nodes + nn);
	return sah_cost(v);
}

int mcpt_host_parse_camera(const char* json_text, mcpt_camera* out) {
	return guarded([&] {
		*out = parse_camera(Json::parse(json_text));
		return 0;
	});
}

int mcpt_host_load_config(const char* path, int configid, mcpt_host_config* out) {
	return guarded([&] {
		RenderConfig c = load_config(path, configid);
		std::memset(out, 0, sizeof(*out));
		auto put = [](char* dst, size_t cap, const std::string& s) {
			std::strncpy(dst, s.c_str(), cap - 1);
			dst[cap - 1] = 0;
		};
		put(out->bvhtype, sizeof(out->bvhtype), c.bvhtype);
		put(out->directory, sizeof(out->directory), c.directory);
		put(out->objname, sizeof(out->objname), c.objname);
		out->width = c.width; out->height = c.height;
		out->maxdepth = c.maxdepth; out->attempt = c.attempt;
		out->opencl = c.opencl; out->testbvh = c.testbvh; out->testall = c.testall;
		out->has_camera = c.has_camera;
		out->res_x = c.res_x; out->res_y = c.res_y;
		if (c.has_camera) out->camera = parse_camera(c.camera);
		if (c.objs.size() > sizeof(out->objs) / sizeof(out->objs[0])) throw std::runtime_error("config: more than 32 models in objname");
		out->nobjs = int(c.objs.size());
		for (size_t i = 0; i < c.objs.size(); ++i) put(out->objs[i], sizeof(out->objs[i]), c.objs[i]);
		return 0;
	});
}

int mcpt_host_write_hdr(const char* path, int w, int h, const float* rgba) {
	return write_hdr(path, w, h, rgba) ? 0 : -1;
}
int mcpt_host_write_png(const char* path, int w, int h, const float* rgba) {
	return write_png(path, w, h, rgba) ? 0 : -1;
}

}  // extern "C"
