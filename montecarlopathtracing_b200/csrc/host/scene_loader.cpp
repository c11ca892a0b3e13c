// OBJ/MTL scene loading -> Triangle[] / Material[] / per-triangle material ids.
//
// Drop-in for ThirdPartyWrapper::loadObject (thirdpartywrapper.cpp:25-99), which drives
// tinyobjloader 2.0.0-rc with triangulate=true.  Contract checked by tests/test_host.py
// against the reference's own code (oracle/_ref): the Triangle and Material arrays are
// memcmp-equal for the shipped scenes and for fuzzed polygon soups.  What has to match:
//   * triangle order = face order in the file; polygons are ear-clipped the way tinyobj does
//     it (projection axes from the first non-degenerate corner, signed area, rotating guess
//     vertex, point-in-triangle rejection) — tiny_obj_loader.h:1359-1580;
//   * material index = position in the .mtl; defaults Ns = 1, Ni = 1 (tiny_obj_loader.h:1297-1298);
//   * classification order and double-precision pre-scales (thirdpartywrapper.cpp:65-97):
//       Ni != 1 -> TRANSPARENT;  Ka > 0 -> LIGHT;  Ns != 1 -> GLOSSY;  else DIFFUSE;
//   * per-triangle geometric normal + material id as the SceneCL constructor adds them
//     (scenebuild.cpp:58-62), using the 4-lane host float math of oclbasic.h:117-146.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <thread>
#ifdef __linux__
#include <sched.h>
#endif

#include "host.hpp"

namespace mcpt::host {
namespace {

constexpr double kHostPi = 3.14159265358;   // oclbasic.h:192

struct MtlRecord {
	std::string name;
	float ambient[3] = {0, 0, 0};
	float diffuse[3] = {0, 0, 0};
	float specular[3] = {0, 0, 0};
	float shininess = 1.0f;
	float ior = 1.0f;
};

inline bool is_space(char c) { return c == ' ' || c == '\t'; }

bool getline_any(std::istream& in, std::string& line) {   // handles \n, \r\n and \r
	line.clear();
	int c;
	bool got = false;
	while ((c = in.get()) != EOF) {
		got = true;
		if (c == '\n') return true;
		if (c == '\r') { if (in.peek() == '\n') in.get(); return true; }
		line.push_back(char(c));
	}
	return got;
}

float parse_float(const char*& p) {
	p += std::strspn(p, " \t");
	char* e = nullptr;
	double v = std::strtod(p, &e);
	if (e == p) { p += std::strcspn(p, " \t"); return 0.0f; }
	p = e;
	return float(v);
}

void parse_float3(const char* p, float out[3]) {
	for (int i = 0; i < 3; ++i) out[i] = parse_float(p);
}

std::vector<MtlRecord> load_mtl(const std::string& path) {
	std::vector<MtlRecord> out;
	std::ifstream in(path, std::ios::binary);
	if (!in) return out;
	std::string line;
	MtlRecord cur;
	bool open = false;
	while (getline_any(in, line)) {
		size_t last = line.find_last_not_of(" \t");
		if (last == std::string::npos) continue;
		line.resize(last + 1);
		const char* t = line.c_str();
		t += std::strspn(t, " \t");
		if (*t == '\0' || *t == '#') continue;
		if (std::strncmp(t, "newmtl", 6) == 0 && is_space(t[6])) {
			if (open && !cur.name.empty()) out.push_back(cur);
			cur = MtlRecord();
			const char* n = t + 7;
			n += std::strspn(n, " \t");
			cur.name = n;
			open = true;
		} else if (t[0] == 'K' && t[1] == 'a' && is_space(t[2])) parse_float3(t + 2, cur.ambient);
		else if (t[0] == 'K' && t[1] == 'd' && is_space(t[2])) parse_float3(t + 2, cur.diffuse);
		else if (t[0] == 'K' && t[1] == 's' && is_space(t[2])) parse_float3(t + 2, cur.specular);
		else if (t[0] == 'N' && t[1] == 'i' && is_space(t[2])) { const char* q = t + 2; cur.ior = parse_float(q); }
		else if (t[0] == 'N' && t[1] == 's' && is_space(t[2])) { const char* q = t + 2; cur.shininess = parse_float(q); }
	}
	if (open && !cur.name.empty()) out.push_back(cur);
	return out;
}

// thirdpartywrapper.cpp:65-97.  `scale * float4` multiplies each lane in double and narrows.
mcpt_material classify(const MtlRecord& m) {
	mcpt_material r{};
	auto scaled = [](double sc, const float c[3]) {
		return mcpt_float4{float(sc * c[0]), float(sc * c[1]), float(sc * c[2]), float(sc * 0.0f)};
	};
	if (m.ior != 1.0f) {
		r.type = MCPT_TRANSPARENT;
		r.Ni = m.ior;
	} else if (m.ambient[0] > 0.0f || m.ambient[1] > 0.0f || m.ambient[2] > 0.0f) {
		r.type = MCPT_LIGHT;
		r.ka_ks = mcpt_float4{m.ambient[0], m.ambient[1], m.ambient[2], 0.0f};
	} else if (m.shininess != 1.0f) {
		r.type = MCPT_GLOSSY;
		r.Ns = m.shininess;
		r.ka_ks = scaled(double(m.shininess + 2) * (2.0 / kHostPi), m.specular);
		r.kd = scaled(1.0 / kHostPi, m.diffuse);
	} else {
		r.type = MCPT_DIFFUSE;
		r.kd = scaled(1.0 / kHostPi, m.diffuse);
	}
	return r;
}

// Even-odd crossing test on a 3-vertex polygon (W. R. Franklin's pnpoly, as tinyobj uses it).
bool inside_tri(const float vx[3], const float vy[3], float tx, float ty) {
	bool c = false;
	for (int i = 0, j = 2; i < 3; j = i++) {
		if (((vy[i] > ty) != (vy[j] > ty)) &&
		    (tx < (vx[j] - vx[i]) * (ty - vy[i]) / (vy[j] - vy[i]) + vx[i]))
			c = !c;
	}
	return c;
}

// Ear clipping with tinyobj's exact decision sequence; emits vertex-index triples.
// `v`/`nv`: the vertex floats read SO FAR (a face may name a vertex that only appears later in the file).
void triangulate_face(const std::vector<int>& face, const float* v, size_t nv, std::vector<int>& out) {
	size_t n = face.size();
	if (n < 3) return;
	auto ok = [&](int vi, size_t off) { return size_t(vi) * 3 + off < nv; };
	// projection axes: drop the dominant axis of the first corner with a non-zero cross product
	size_t ax0 = 1, ax1 = 2;
	for (size_t k = 0; k < n; ++k) {
		int a = face[k % n], b = face[(k + 1) % n], c = face[(k + 2) % n];
		if (!ok(a, 2) || !ok(b, 2) || !ok(c, 2)) continue;
		float e0x = v[b * 3 + 0] - v[a * 3 + 0], e0y = v[b * 3 + 1] - v[a * 3 + 1], e0z = v[b * 3 + 2] - v[a * 3 + 2];
		float e1x = v[c * 3 + 0] - v[b * 3 + 0], e1y = v[c * 3 + 1] - v[b * 3 + 1], e1z = v[c * 3 + 2] - v[b * 3 + 2];
		float cx = std::fabs(e0y * e1z - e0z * e1y);
		float cy = std::fabs(e0z * e1x - e0x * e1z);
		float cz = std::fabs(e0x * e1y - e0y * e1x);
		const float eps = std::numeric_limits<float>::epsilon();
		if (cx > eps || cy > eps || cz > eps) {
			if (!(cx > cy && cx > cz)) {
				ax0 = 0;
				if (cz > cx && cz > cy) ax1 = 1;
			}
			break;
		}
	}
	float area = 0;
	for (size_t k = 0; k < n; ++k) {
		int a = face[k % n], b = face[(k + 1) % n];
		if (!ok(a, ax0) || !ok(a, ax1) || !ok(b, ax0) || !ok(b, ax1)) continue;
		float ax = v[a * 3 + ax0], ay = v[a * 3 + ax1], bx = v[b * 3 + ax0], by = v[b * 3 + ax1];
		area += (ax * by - ay * bx) * 0.5f;
	}
	std::vector<int> rem = face;
	size_t guess = 0, iters_left = n, prev_count = rem.size();
	while (rem.size() > 3 && iters_left > 0) {
		size_t m = rem.size();
		if (guess >= m) guess -= m;
		if (prev_count != m) { prev_count = m; iters_left = m; }
		else --iters_left;
		int ind[3];
		float vx[3], vy[3];
		for (int k = 0; k < 3; ++k) {
			ind[k] = rem[(guess + k) % m];
			if (!ok(ind[k], ax0) || !ok(ind[k], ax1)) { vx[k] = vy[k] = 0.0f; }
			else { vx[k] = v[ind[k] * 3 + ax0]; vy[k] = v[ind[k] * 3 + ax1]; }
		}
		float e0x = vx[1] - vx[0], e0y = vy[1] - vy[0], e1x = vx[2] - vx[1], e1y = vy[2] - vy[1];
		float cr = e0x * e1y - e0y * e1x;
		if (cr * area < 0.0f) { ++guess; continue; }       // reflex corner
		bool overlap = false;
		for (size_t o = 3; o < m; ++o) {
			size_t idx = (guess + o) % m;
			int ov = rem[idx];
			if (!ok(ov, ax0) || !ok(ov, ax1)) continue;
			if (inside_tri(vx, vy, v[ov * 3 + ax0], v[ov * 3 + ax1])) { overlap = true; break; }
		}
		if (overlap) { ++guess; continue; }
		out.push_back(ind[0]); out.push_back(ind[1]); out.push_back(ind[2]);
		rem.erase(rem.begin() + long((guess + 1) % m));
	}
	if (rem.size() == 3) { out.push_back(rem[0]); out.push_back(rem[1]); out.push_back(rem[2]); }
}

inline mcpt_float4 sub4(const mcpt_float4& a, const mcpt_float4& b) { return {a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }

}  // namespace

Scene load_object(const std::string& directory, const std::string& objname, bool with_normals) {
	Scene sc;
	std::ifstream in(directory + objname, std::ios::binary);
	if (!in) throw std::runtime_error("cannot open " + directory + objname);
	std::vector<float> verts;
	std::vector<MtlRecord> mtl;
	std::map<std::string, int> mtl_index;
	int cur_mat = -1;
	std::string line;
	std::vector<int> face, tri_idx;
	while (getline_any(in, line)) {
		size_t last = line.find_last_not_of(" \t");
		if (last == std::string::npos) continue;
		line.resize(last + 1);
		const char* t = line.c_str();
		t += std::strspn(t, " \t");
		if (*t == '\0' || *t == '#') continue;
		if (t[0] == 'v' && is_space(t[1])) {
			const char* p = t + 2;
			for (int i = 0; i < 3; ++i) verts.push_back(parse_float(p));
		} else if (t[0] == 'f' && is_space(t[1])) {
			const char* p = t + 2;
			face.clear();
			int nverts_now = int(verts.size() / 3);
			for (;;) {
				p += std::strspn(p, " \t");
				if (*p == '\0') break;
				char* e = nullptr;
				long i = std::strtol(p, &e, 10);
				if (e == p) break;
				int vi = i > 0 ? int(i - 1) : (i < 0 ? nverts_now + int(i) : 0);
				face.push_back(vi);
				p = e;
				p += std::strcspn(p, " \t");    // skip /vt/vn
			}
			tri_idx.clear();
			triangulate_face(face, verts.data(), verts.size(), tri_idx);
			for (size_t k = 0; k + 2 < tri_idx.size() + 0 && k < tri_idx.size(); k += 3) {
				mcpt_triangle tr{};
				for (int c = 0; c < 3; ++c) {
					size_t b = size_t(tri_idx[k + c]) * 3;
					if (b + 2 >= verts.size()) throw std::runtime_error("obj: vertex index out of range");
					tr.v[c] = mcpt_float4{verts[b], verts[b + 1], verts[b + 2], 0.0f};
				}
				sc.triangles.push_back(tr);
				sc.material_ids.push_back(cur_mat);
			}
		} else if (std::strncmp(t, "usemtl", 6) == 0 && is_space(t[6])) {
			std::string name = t + 7;
			auto it = mtl_index.find(name);
			cur_mat = it == mtl_index.end() ? -1 : it->second;
		} else if (std::strncmp(t, "mtllib", 6) == 0 && is_space(t[6])) {
			std::string rest = t + 7;
			size_t pos = 0;
			while (pos < rest.size()) {     // first file that loads wins (tiny_obj_loader.h:2346-2375)
				size_t sp = rest.find(' ', pos);
				std::string fn = rest.substr(pos, sp == std::string::npos ? std::string::npos : sp - pos);
				pos = sp == std::string::npos ? rest.size() : sp + 1;
				if (fn.empty()) continue;
				auto recs = load_mtl(directory + fn);
				if (recs.empty()) continue;
				for (auto& r : recs) {
					mtl_index.emplace(r.name, int(mtl.size()));
					mtl.push_back(r);
				}
				break;
			}
		}
	}
	for (auto& m : mtl) {
		sc.materials.push_back(classify(m));
		sc.material_names.push_back(m.name);
	}
	if (with_normals) {
		for (size_t i = 0; i < sc.triangles.size(); ++i) {
			auto& tr = sc.triangles[i];
			mcpt_float4 a = sub4(tr.v[1], tr.v[0]), b = sub4(tr.v[2], tr.v[0]);
			mcpt_float4 n{0, 0, 0, 0};                       // oclbasic.h:117-125
			n.x = a.y * b.z - a.z * b.y;
			n.y = -a.x * b.z + a.z * b.x;
			n.z = a.x * b.y - a.y * b.x;
			float s = 0;                                     // 4-lane dot, oclbasic.h:127-135
			s += n.x * n.x; s += n.y * n.y; s += n.z * n.z; s += n.w * n.w;
			s = std::sqrt(s);
			tr.nx = n.x / s; tr.ny = n.y / s; tr.nz = n.z / s;
			tr.material_id = sc.material_ids[i];
		}
	}
	return sc;
}

// ------------------------------------------------------------------------------------------------------
// Parallel loader: same bytes as load_object above, which stays the pinned definition.
// The file is read once and cut at line ends into one chunk per thread.  Pass 1 (parallel) turns each chunk
// into vertex floats, raw face indices and the few state-changing lines (usemtl / mtllib) with their position;
// a short serial pass replays those state changes in file order (material table, current material) and turns
// per-chunk vertex counts into offsets; pass 2 (parallel) resolves indices against "vertices read so far",
// ear-clips with the same routine and writes triangles, material ids and normals in place.
namespace {

struct ChunkEvent {
	size_t face;              // number of faces of this chunk read before the line
	bool is_lib;              // mtllib (true) or usemtl (false)
	std::string arg;
};
struct Chunk {
	const char* begin = nullptr;
	const char* end = nullptr;
	std::vector<float> verts;
	std::vector<long> raw;            // face indices as written (1-based, negative = relative)
	std::vector<uint32_t> face_len;
	std::vector<uint32_t> face_nv;    // vertices of THIS chunk read before the face
	std::vector<ChunkEvent> events;
	// after the serial pass
	size_t vert_base = 0;             // vertices (not floats) in earlier chunks
	std::vector<std::pair<size_t, int>> mat_runs;   // (first face, material) in face order
	std::vector<mcpt_triangle> tris;
	std::vector<int> tri_mat;
	std::string error;
};

// one line [b, e) without its terminator; p advances past \n, \r\n or \r
inline bool next_line(const char*& p, const char* end, const char*& b, const char*& e) {
	if (p >= end) return false;
	b = p;
	while (p < end && *p != '\n' && *p != '\r') ++p;
	e = p;
	if (p < end) {
		if (*p == '\r' && p + 1 < end && p[1] == '\n') p += 2;
		else ++p;
	}
	return true;
}

void parse_chunk(Chunk& c) {
	const char* p = c.begin;
	const char *b, *e;
	std::string line;
	while (next_line(p, c.end, b, e)) {
		while (e > b && is_space(e[-1])) --e;
		if (e == b) continue;
		line.assign(b, e);
		const char* t = line.c_str();
		t += std::strspn(t, " \t");
		if (*t == '\0' || *t == '#') continue;
		if (t[0] == 'v' && is_space(t[1])) {
			const char* q = t + 2;
			for (int i = 0; i < 3; ++i) c.verts.push_back(parse_float(q));
		} else if (t[0] == 'f' && is_space(t[1])) {
			const char* q = t + 2;
			uint32_t len = 0;
			for (;;) {
				q += std::strspn(q, " \t");
				if (*q == '\0') break;
				char* ep = nullptr;
				long i = std::strtol(q, &ep, 10);
				if (ep == q) break;
				c.raw.push_back(i);
				++len;
				q = ep;
				q += std::strcspn(q, " \t");
			}
			c.face_len.push_back(len);
			c.face_nv.push_back(uint32_t(c.verts.size() / 3));
		} else if (std::strncmp(t, "usemtl", 6) == 0 && is_space(t[6])) {
			c.events.push_back({c.face_len.size(), false, std::string(t + 7)});
		} else if (std::strncmp(t, "mtllib", 6) == 0 && is_space(t[6])) {
			c.events.push_back({c.face_len.size(), true, std::string(t + 7)});
		}
	}
}

void emit_chunk(Chunk& c, const std::vector<float>& verts) {
	std::vector<int> face, tri_idx;
	size_t r = 0, run = 0;
	int cur_mat = c.mat_runs.empty() ? -1 : c.mat_runs[0].second;
	for (size_t f = 0; f < c.face_len.size(); ++f) {
		while (run + 1 < c.mat_runs.size() && c.mat_runs[run + 1].first <= f) ++run;
		if (!c.mat_runs.empty()) cur_mat = c.mat_runs[run].second;
		const int nverts_now = int(c.vert_base + c.face_nv[f]);
		const size_t nfloats_now = size_t(nverts_now) * 3;
		face.clear();
		for (uint32_t k = 0; k < c.face_len[f]; ++k) {
			const long i = c.raw[r++];
			face.push_back(i > 0 ? int(i - 1) : (i < 0 ? nverts_now + int(i) : 0));
		}
		tri_idx.clear();
		triangulate_face(face, verts.data(), nfloats_now, tri_idx);
		for (size_t k = 0; k + 2 < tri_idx.size(); k += 3) {
			mcpt_triangle tr{};
			for (int v = 0; v < 3; ++v) {
				size_t b = size_t(tri_idx[k + v]) * 3;
				if (b + 2 >= nfloats_now) { c.error = "obj: vertex index out of range"; return; }
				tr.v[v] = mcpt_float4{verts[b], verts[b + 1], verts[b + 2], 0.0f};
			}
			c.tris.push_back(tr);
			c.tri_mat.push_back(cur_mat);
		}
	}
}

template <class F>
void run_threads(size_t n, F&& fn) {
	std::vector<std::thread> th;
	for (size_t i = 1; i < n; ++i) th.emplace_back([&fn, i] { fn(i); });
	fn(0);
	for (auto& t : th) t.join();
}

}  // namespace

Scene load_object_parallel(const std::string& directory, const std::string& objname, bool with_normals, int threads) {
	if (threads <= 0) {
		threads = int(std::thread::hardware_concurrency());
#ifdef __linux__
		cpu_set_t set;                                    // a container may be allowed fewer cores than the machine has
		if (sched_getaffinity(0, sizeof(set), &set) == 0) threads = std::min(threads, std::max(1, CPU_COUNT(&set)));
#endif
		threads = std::min(threads, 32);
	}
	if (threads <= 1) return load_object(directory, objname, with_normals);
	std::string data;
	{
		std::ifstream in(directory + objname, std::ios::binary);
		if (!in) throw std::runtime_error("cannot open " + directory + objname);
		in.seekg(0, std::ios::end);
		const std::streamoff sz = in.tellg();
		in.seekg(0, std::ios::beg);
		data.resize(size_t(sz > 0 ? sz : 0));
		if (!data.empty()) in.read(&data[0], std::streamsize(data.size()));
	}
	const char* base = data.data();
	const char* end = base + data.size();
	size_t nchunks = std::min<size_t>(size_t(threads), std::max<size_t>(1, data.size() / 65536));
	std::vector<Chunk> chunks(nchunks);
	{
		const char* p = base;
		for (size_t i = 0; i < nchunks; ++i) {
			chunks[i].begin = p;
			const char* target = (i + 1 == nchunks) ? end : base + data.size() * (i + 1) / nchunks;
			if (target < p) target = p;
			const char *b, *e;
			const char* q = target;
			if (q < end) next_line(q, end, b, e);      // finish the line `target` falls into
			chunks[i].end = p = q;
		}
		chunks.back().end = end;
	}
	run_threads(nchunks, [&](size_t i) { parse_chunk(chunks[i]); });

	// ---- serial replay of the state-changing lines, vertex offsets
	Scene sc;
	std::vector<MtlRecord> mtl;
	std::map<std::string, int> mtl_index;
	int cur_mat = -1;
	size_t nverts = 0;
	for (Chunk& c : chunks) {
		c.vert_base = nverts;
		nverts += c.verts.size() / 3;
		c.mat_runs.push_back({0, cur_mat});
		for (const ChunkEvent& ev : c.events) {
			if (!ev.is_lib) {
				auto it = mtl_index.find(ev.arg);
				cur_mat = it == mtl_index.end() ? -1 : it->second;
				c.mat_runs.push_back({ev.face, cur_mat});
				continue;
			}
			size_t pos = 0;
			while (pos < ev.arg.size()) {     // first file that loads wins (tiny_obj_loader.h:2346-2375)
				size_t sp = ev.arg.find(' ', pos);
				std::string fn = ev.arg.substr(pos, sp == std::string::npos ? std::string::npos : sp - pos);
				pos = sp == std::string::npos ? ev.arg.size() : sp + 1;
				if (fn.empty()) continue;
				auto recs = load_mtl(directory + fn);
				if (recs.empty()) continue;
				for (auto& r : recs) {
					mtl_index.emplace(r.name, int(mtl.size()));
					mtl.push_back(r);
				}
				break;
			}
		}
	}
	std::vector<float> verts(nverts * 3);
	run_threads(nchunks, [&](size_t i) {
		if (!chunks[i].verts.empty())
			std::memcpy(verts.data() + chunks[i].vert_base * 3, chunks[i].verts.data(), chunks[i].verts.size() * sizeof(float));
	});
	run_threads(nchunks, [&](size_t i) { emit_chunk(chunks[i], verts); });
	size_t ntris = 0;
	std::vector<size_t> tri_base(nchunks);
	for (size_t i = 0; i < nchunks; ++i) {
		// the serial loader stops at the first bad face; everything before it is irrelevant once it throws
		if (!chunks[i].error.empty()) throw std::runtime_error(chunks[i].error);
		tri_base[i] = ntris;
		ntris += chunks[i].tris.size();
	}
	sc.triangles.resize(ntris);
	sc.material_ids.resize(ntris);
	run_threads(nchunks, [&](size_t i) {
		Chunk& c = chunks[i];
		for (size_t k = 0; k < c.tris.size(); ++k) {
			mcpt_triangle tr = c.tris[k];
			if (with_normals) {
				mcpt_float4 a = sub4(tr.v[1], tr.v[0]), b = sub4(tr.v[2], tr.v[0]);
				mcpt_float4 n{0, 0, 0, 0};                       // oclbasic.h:117-125
				n.x = a.y * b.z - a.z * b.y;
				n.y = -a.x * b.z + a.z * b.x;
				n.z = a.x * b.y - a.y * b.x;
				float s = 0;                                     // 4-lane dot, oclbasic.h:127-135
				s += n.x * n.x; s += n.y * n.y; s += n.z * n.z; s += n.w * n.w;
				s = std::sqrt(s);
				tr.nx = n.x / s; tr.ny = n.y / s; tr.nz = n.z / s;
				tr.material_id = c.tri_mat[k];
			}
			sc.triangles[tri_base[i] + k] = tr;
			sc.material_ids[tri_base[i] + k] = c.tri_mat[k];
		}
	});
	for (auto& m : mtl) {
		sc.materials.push_back(classify(m));
		sc.material_names.push_back(m.name);
	}
	return sc;
}

}  // namespace mcpt::host
