// Binary scene cache.  The reference re-parses the OBJ through tinyobj and rebuilds the BVH on every
// start (thirdpartywrapper.cpp:25-99, tiny_obj_loader.h:2063, scenebuild.cpp:50-101); here the results of that work -
// Triangle[], Material[], the per-triangle material ids, the BVHNode[] of the configured "bvhtype" and, opaque to this
// file, the traversal layout libmcpt_cuda derives from them - are kept in one versioned file keyed by the CONTENT of
// the OBJ and of every MTL file it names.  A cache hit returns the same bytes a fresh load produces (tests/test_scene_cache.py).
//
//   file = Header | Triangle[ntris] | Material[nmats] | int32 ids[ntris] | BVHNode[nnodes] | derived[derived_bytes]
//
// Written to a temporary name and renamed, so concurrent writers (one process per GPU) never expose a partial file.
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

#include <unistd.h>

#include "../common/hash128.hpp"
#include "host.hpp"

namespace mcpt::host {
namespace {

constexpr char MAGIC[8] = {'M', 'C', 'P', 'T', 'S', 'C', 'N', '1'};
constexpr uint32_t VERSION = 1;

struct Header {
	char magic[8];
	uint32_t version;
	uint32_t struct_sizes;      // sizeof(Triangle) | sizeof(Material) << 8 | sizeof(BVHNode) << 16: layout drift invalidates
	uint64_t key[2];
	int32_t ntris, nmats, nnodes, with_normals;
	uint64_t derived_bytes;
	uint64_t payload_hash[2];   // over the five sections in file order
};
static_assert(sizeof(Header) == 72, "scene cache header");

constexpr uint32_t struct_sizes() {
	return uint32_t(sizeof(mcpt_triangle)) | uint32_t(sizeof(mcpt_material)) << 8 | uint32_t(sizeof(mcpt_bvh_node)) << 16;
}

bool read_file(const std::string& path, std::vector<char>& out) {
	std::FILE* f = std::fopen(path.c_str(), "rb");
	if (!f) return false;
	std::fseek(f, 0, SEEK_END);
	long n = std::ftell(f);
	std::fseek(f, 0, SEEK_SET);
	out.resize(n > 0 ? size_t(n) : 0);
	const bool ok = out.empty() || std::fread(out.data(), 1, out.size(), f) == out.size();
	std::fclose(f);
	return ok;
}

struct File {
	std::FILE* f = nullptr;
	~File() { if (f) std::fclose(f); }
};

}  // namespace

SceneKey scene_key(const std::string& directory, const std::string& objname, const std::string& bvhtype, bool with_normals) {
	std::vector<char> obj;
	if (!read_file(directory + objname, obj)) throw std::runtime_error("cannot open " + directory + objname);
	Hash128 h;
	const uint32_t tag[4] = {VERSION, struct_sizes(), with_normals ? 1u : 0u, 0u};
	hash_update(h, tag, sizeof(tag), 1);
	hash_update(h, bvhtype.c_str());
	hash_update(h, obj.data(), obj.size());
	// every file named on a mtllib line, in order of appearance (the loader takes the first one of a line that
	// loads, thirdpartywrapper.cpp via tiny_obj_loader.h:2346-2375: hashing all of them is a superset)
	const char* p = obj.data();
	const char* end = p + obj.size();
	while (p < end) {
		const char* nl = static_cast<const char*>(std::memchr(p, '\n', size_t(end - p)));
		const char* e = nl ? nl : end;
		const char* t = p;
		while (t < e && (*t == ' ' || *t == '\t')) ++t;
		if (e - t > 7 && std::memcmp(t, "mtllib", 6) == 0 && (t[6] == ' ' || t[6] == '\t')) {
			std::string rest(t + 7, e);
			while (!rest.empty() && (rest.back() == '\r' || rest.back() == ' ' || rest.back() == '\t')) rest.pop_back();
			size_t pos = 0;
			while (pos < rest.size()) {
				size_t sp = rest.find(' ', pos);
				std::string fn = rest.substr(pos, sp == std::string::npos ? std::string::npos : sp - pos);
				pos = sp == std::string::npos ? rest.size() : sp + 1;
				if (fn.empty()) continue;
				hash_update(h, fn.c_str());
				std::vector<char> mtl;
				if (read_file(directory + fn, mtl)) hash_update(h, mtl.data(), mtl.size());
				else hash_update(h, "<absent>");
			}
		}
		p = nl ? nl + 1 : end;
	}
	return SceneKey{{h.a, h.b}};
}

std::string scene_cache_path(const std::string& directory, const std::string& objname, const std::string& bvhtype) {
	// next to the OBJ unless MCPT_CACHE_DIR names another place (read-only scene directories)
	const char* dir = std::getenv("MCPT_CACHE_DIR");
	std::string base = dir && *dir ? std::string(dir) + "/" : directory;
	std::string name = objname;
	if (dir && *dir) for (char& c : name) if (c == '/') c = '_';
	return base + name + "." + bvhtype + ".mcptcache";
}

void scene_cache_store(const std::string& path, const SceneKey& key, const CachedScene& s) {
	if (s.material_ids.size() != s.triangles.size()) throw std::runtime_error("scene cache: material id count != triangle count");
	Header h{};
	std::memcpy(h.magic, MAGIC, 8);
	h.version = VERSION;
	h.struct_sizes = struct_sizes();
	h.key[0] = key.h[0]; h.key[1] = key.h[1];
	h.ntris = int32_t(s.triangles.size());
	h.nmats = int32_t(s.materials.size());
	h.nnodes = int32_t(s.nodes.size());
	h.with_normals = s.with_normals ? 1 : 0;
	h.derived_bytes = s.derived.size();
	Hash128 ph;
	hash_update(ph, s.triangles.data(), s.triangles.size() * sizeof(mcpt_triangle));
	hash_update(ph, s.materials.data(), s.materials.size() * sizeof(mcpt_material));
	hash_update(ph, s.material_ids.data(), s.material_ids.size() * sizeof(int));
	hash_update(ph, s.nodes.data(), s.nodes.size() * sizeof(mcpt_bvh_node));
	hash_update(ph, s.derived.data(), s.derived.size());
	h.payload_hash[0] = ph.a; h.payload_hash[1] = ph.b;
	const std::string tmp = path + ".tmp." + std::to_string(long(getpid()));
	{
		File out;
		out.f = std::fopen(tmp.c_str(), "wb");
		if (!out.f) throw std::runtime_error("scene cache: cannot write " + tmp + ": " + std::strerror(errno));
		auto put = [&](const void* p, size_t n) {
			if (n && std::fwrite(p, 1, n, out.f) != n) throw std::runtime_error("scene cache: short write to " + tmp);
		};
		put(&h, sizeof(h));
		put(s.triangles.data(), s.triangles.size() * sizeof(mcpt_triangle));
		put(s.materials.data(), s.materials.size() * sizeof(mcpt_material));
		put(s.material_ids.data(), s.material_ids.size() * sizeof(int));
		put(s.nodes.data(), s.nodes.size() * sizeof(mcpt_bvh_node));
		put(s.derived.data(), s.derived.size());
		if (std::fflush(out.f) != 0) throw std::runtime_error("scene cache: flush failed for " + tmp);
	}
	if (std::rename(tmp.c_str(), path.c_str()) != 0) {
		std::remove(tmp.c_str());
		throw std::runtime_error("scene cache: cannot rename to " + path + ": " + std::strerror(errno));
	}
}

bool scene_cache_load_raw(const std::string& path, const SceneKey& key, RawScene& s, std::string* why) {
	auto miss = [&](const std::string& m) { if (why) *why = m; return false; };
	File in;
	in.f = std::fopen(path.c_str(), "rb");
	if (!in.f) return miss("no cache file");
	Header h{};
	if (std::fread(&h, 1, sizeof(h), in.f) != sizeof(h)) return miss("truncated header");
	if (std::memcmp(h.magic, MAGIC, 8) != 0) return miss("not a scene cache file");
	if (h.version != VERSION || h.struct_sizes != struct_sizes()) return miss("other format version");
	if (h.key[0] != key.h[0] || h.key[1] != key.h[1]) return miss("stale: OBJ / MTL content or bvhtype changed");
	if (h.ntris <= 0 || h.nmats < 0 || h.nnodes < 0 || h.derived_bytes > (uint64_t(1) << 40)) return miss("implausible counts");
	std::fseek(in.f, 0, SEEK_END);
	const uint64_t size = uint64_t(std::ftell(in.f));
	const uint64_t want = sizeof(Header) + uint64_t(h.ntris) * (sizeof(mcpt_triangle) + sizeof(int)) +
	                      uint64_t(h.nmats) * sizeof(mcpt_material) + uint64_t(h.nnodes) * sizeof(mcpt_bvh_node) + h.derived_bytes;
	if (size != want) return miss("file size does not match its header");
	std::fseek(in.f, long(sizeof(Header)), SEEK_SET);
	// every section is read straight into the buffer the caller will own (mcpt_host_free): one copy of 200 MB, not three
	const size_t bytes[5] = {size_t(h.ntris) * sizeof(mcpt_triangle), size_t(h.nmats) * sizeof(mcpt_material),
	                         size_t(h.ntris) * sizeof(int), size_t(h.nnodes) * sizeof(mcpt_bvh_node), size_t(h.derived_bytes)};
	void* buf[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
	auto release = [&] { for (void*& b : buf) { std::free(b); b = nullptr; } };
	Hash128 ph;
	for (int k = 0; k < 5; ++k) {
		buf[k] = std::malloc(bytes[k] + 1);
		if (!buf[k]) { release(); return miss("out of memory"); }
		if (bytes[k] && std::fread(buf[k], 1, bytes[k], in.f) != bytes[k]) { release(); return miss("short read"); }
		hash_update(ph, buf[k], bytes[k]);
	}
	if (ph.a != h.payload_hash[0] || ph.b != h.payload_hash[1]) { release(); return miss("payload checksum mismatch (corrupt file)"); }
	s.triangles = static_cast<mcpt_triangle*>(buf[0]);
	s.materials = static_cast<mcpt_material*>(buf[1]);
	s.material_ids = static_cast<int*>(buf[2]);
	s.nodes = static_cast<mcpt_bvh_node*>(buf[3]);
	s.derived = static_cast<unsigned char*>(buf[4]);
	s.ntris = h.ntris; s.nmats = h.nmats; s.nnodes = h.nnodes; s.derived_bytes = size_t(h.derived_bytes);
	s.with_normals = h.with_normals != 0;
	return true;
}

bool scene_cache_load(const std::string& path, const SceneKey& key, CachedScene& s, std::string* why) {
	RawScene r{};
	if (!scene_cache_load_raw(path, key, r, why)) return false;
	s.with_normals = r.with_normals;
	s.triangles.assign(r.triangles, r.triangles + r.ntris);
	s.materials.assign(r.materials, r.materials + r.nmats);
	s.material_ids.assign(r.material_ids, r.material_ids + r.ntris);
	s.nodes.assign(r.nodes, r.nodes + r.nnodes);
	s.derived.assign(r.derived, r.derived + r.derived_bytes);
	std::free(r.triangles); std::free(r.materials); std::free(r.material_ids); std::free(r.nodes); std::free(r.derived);
	return true;
}

}  // namespace mcpt::host
