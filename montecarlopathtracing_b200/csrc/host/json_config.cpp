// JSON(+'#' comments) reader, config.json loader and camera basis.
// Reference behaviour: config.cpp:70-125 (field set, defaults), json.hpp:3037-3043 ('#' comments),
// auxiliary.cpp:20-71 (parseCamera).
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <sstream>

#include "host.hpp"

namespace mcpt::host {

// ---------------------------------------------------------------------------------- Json
namespace {
struct Parser {
	const std::string& s;
	size_t p = 0;
	explicit Parser(const std::string& text) : s(text) {}

	[[noreturn]] void fail(const char* what) const {
		throw std::runtime_error(std::string("json: ") + what + " at offset " + std::to_string(p));
	}
	void skip() {
		for (;;) {
			while (p < s.size() && (s[p] == ' ' || s[p] == '\t' || s[p] == '\n' || s[p] == '\r')) ++p;
			if (p < s.size() && s[p] == '#') {           // comment runs to end of line
				while (p < s.size() && s[p] != '\n') ++p;
				continue;
			}
			return;
		}
	}
	bool eat(char c) {
		skip();
		if (p < s.size() && s[p] == c) { ++p; return true; }
		return false;
	}
	Json value() {
		skip();
		if (p >= s.size()) fail("unexpected end");
		char c = s[p];
		Json j;
		if (c == '{') {
			++p;
			j.kind = Json::Object;
			if (eat('}')) return j;
			do {
				skip();
				if (p < s.size() && s[p] == '}') break;     // trailing comma (the shipped config.json has one once its '#' lines are dropped)
				if (p >= s.size() || s[p] != '"') fail("expected key string");
				std::string k = string_body();
				if (!eat(':')) fail("expected ':'");
				j.obj.emplace_back(std::move(k), value());
			} while (eat(','));
			if (!eat('}')) fail("expected '}'");
		} else if (c == '[') {
			++p;
			j.kind = Json::Array;
			if (eat(']')) return j;
			do {
				skip();
				if (p < s.size() && s[p] == ']') break;     // trailing comma
				j.arr.push_back(value());
			} while (eat(','));
			if (!eat(']')) fail("expected ']'");
		} else if (c == '"') {
			j.kind = Json::String;
			j.str = string_body();
		} else if (s.compare(p, 4, "true") == 0) {
			p += 4; j.kind = Json::Bool; j.b = true;
		} else if (s.compare(p, 5, "false") == 0) {
			p += 5; j.kind = Json::Bool; j.b = false;
		} else if (s.compare(p, 4, "null") == 0) {
			p += 4;
		} else {
			const char* b = s.c_str() + p;
			char* e = nullptr;
			j.num = std::strtod(b, &e);
			if (e == b) fail("unexpected character");
			p += size_t(e - b);
			j.kind = Json::Number;
		}
		return j;
	}
	std::string string_body() {
		++p;  // opening quote
		std::string out;
		while (p < s.size() && s[p] != '"') {
			char c = s[p++];
			if (c == '\\') {
				if (p >= s.size()) fail("bad escape");
				char e = s[p++];
				switch (e) {
				case 'n': out += '\n'; break;
				case 't': out += '\t'; break;
				case 'r': out += '\r'; break;
				case 'b': out += '\b'; break;
				case 'f': out += '\f'; break;
				case 'u': {
					if (p + 4 > s.size()) fail("bad \\u escape");
					unsigned cp = unsigned(std::strtoul(s.substr(p, 4).c_str(), nullptr, 16));
					p += 4;
					if (cp < 0x80) out += char(cp);
					else if (cp < 0x800) { out += char(0xC0 | (cp >> 6)); out += char(0x80 | (cp & 0x3F)); }
					else { out += char(0xE0 | (cp >> 12)); out += char(0x80 | ((cp >> 6) & 0x3F)); out += char(0x80 | (cp & 0x3F)); }
					break;
				}
				default: out += e;   // \" \\ \/
				}
			} else {
				out += c;
			}
		}
		if (p >= s.size()) fail("unterminated string");
		++p;  // closing quote
		return out;
	}
};
}  // namespace

Json Json::parse(const std::string& text) {
	Parser ps(text);
	Json j = ps.value();
	ps.skip();
	if (ps.p != text.size()) ps.fail("trailing characters");
	return j;
}
bool Json::has(const std::string& key) const {
	for (auto& kv : obj) if (kv.first == key) return true;
	return false;
}
const Json& Json::at(const std::string& key) const {
	for (auto& kv : obj) if (kv.first == key) return kv.second;
	throw std::runtime_error("json: missing key '" + key + "'");
}
const Json& Json::at(size_t i) const {
	if (kind != Array || i >= arr.size()) throw std::runtime_error("json: array index out of range");
	return arr[i];
}
double Json::number() const {
	if (kind != Number) throw std::runtime_error("json: not a number");
	return num;
}
const std::string& Json::string() const {
	if (kind != String) throw std::runtime_error("json: not a string");
	return str;
}
bool Json::boolean() const {
	if (kind == Bool) return b;
	if (kind == Number) return num != 0.0;
	throw std::runtime_error("json: not a boolean");
}

// ---------------------------------------------------------------------------------- config
RenderConfig config_from_json(const Json& root, int configid_override) {
	const Json& configs = root.at("config");
	int id = configid_override >= 0 ? configid_override : int(root.at("configid").number());
	const Json& c = configs.at(size_t(id));
	RenderConfig r;
	if (c.has("bvhtype") && !c.at("bvhtype").string().empty()) r.bvhtype = c.at("bvhtype").string();
	r.testall = c.has("testall") && c.at("testall").boolean();
	r.directory = c.at("directory").string();
	if (r.testall) {                               // objname is an array in this mode (config.cpp:91-95)
		const Json& names = c.at("objname");
		for (size_t i = 0; i < names.size(); ++i) r.objs.push_back(names.at(i).string());
		return r;
	}
	if (c.has("camera")) { r.camera = c.at("camera"); r.has_camera = r.camera.kind == Json::Object; }
	r.objname = c.at("objname").string();
	r.width = int(c.at("width").number());
	r.height = int(c.at("height").number());
	if (r.has_camera && r.camera.has("resolution")) {
		r.res_x = int(r.camera.at("resolution").at(0).number());
		r.res_y = int(r.camera.at("resolution").at(1).number());
	}
	r.testbvh = c.has("testbvh") && c.at("testbvh").boolean();
	if (r.testbvh) return r;
	if (c.has("opencl")) r.opencl = c.at("opencl").boolean();
	r.maxdepth = int(c.at("maxdepth").number());
	r.attempt = int(c.at("attempt").number());
	return r;
}

RenderConfig load_config(const std::string& path, int configid_override) {
	std::ifstream fin(path);
	if (!fin) throw std::runtime_error("cannot open " + path);
	std::stringstream ss;
	ss << fin.rdbuf();
	return config_from_json(Json::parse(ss.str()), configid_override);
}

// ---------------------------------------------------------------------------------- camera
namespace {
constexpr double kHostPi = 3.14159265358;          // oclbasic.h:192 (NOT the libm constant)

inline mcpt_float4 cross4(const mcpt_float4& a, const mcpt_float4& b) {   // oclbasic.h:117-125
	mcpt_float4 r{0, 0, 0, 0};
	r.x = a.y * b.z - a.z * b.y;
	r.y = -a.x * b.z + a.z * b.x;
	r.z = a.x * b.y - a.y * b.x;
	return r;
}
inline float dot4(const mcpt_float4& a, const mcpt_float4& b) {           // oclbasic.h:127-135
	float s = 0;
	s += a.x * b.x; s += a.y * b.y; s += a.z * b.z; s += a.w * b.w;
	return s;
}
inline mcpt_float4 normalize4(const mcpt_float4& a) {                      // oclbasic.h:137-146
	float l = std::sqrt(dot4(a, a));
	return mcpt_float4{a.x / l, a.y / l, a.z / l, a.w / l};
}
}  // namespace

mcpt_camera parse_camera(const Json& cam) {
	mcpt_camera c{};
	c.camera_type = 0;
	const Json& pos = cam.at("position");
	const Json& look = cam.at("lookat");
	const Json& up = cam.at("up");
	float* ce = &c.center.x;
	for (int i = 0; i < 3; ++i) ce[i] = float(pos.at(size_t(i)).number());
	mcpt_float4 lookat{};
	float* la = &lookat.x;
	for (int i = 0; i < 3; ++i) la[i] = float(look.at(size_t(i)).number());
	// `float4 lookat;` is uninitialised in the reference, so lookat.w - center.w is garbage there;
	// the next statement overwrites direction.w with 0, so the garbage never escapes.
	c.direction = mcpt_float4{lookat.x - c.center.x, lookat.y - c.center.y, lookat.z - c.center.z, 0.0f};
	float* u = &c.up.x;
	for (int i = 0; i < 3; ++i) u[i] = float(up.at(size_t(i)).number());
	// float * double / float -> evaluated in double, narrowed once (auxiliary.cpp:42)
	c.arg = float(double(float(cam.at("fov").number())) * kHostPi / double(180.0f));
	mcpt_float4 horizontal = cross4(c.direction, c.up);
	c.up = cross4(horizontal, c.direction);
	c.tmin = 0;
	c.direction = normalize4(c.direction);
	c.up = normalize4(c.up);
	c.horizontal = normalize4(horizontal);
	return c;
}

}  // namespace mcpt::host
