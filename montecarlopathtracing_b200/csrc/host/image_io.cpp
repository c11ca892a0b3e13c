// Image output: Radiance .hdr (the reference's only output, thirdpartywrapper.cpp:14-23 via
// stb_image_write 1.13) and an 8-bit PNG of the display transform (testkernel.cl:9-14).
//
// The .hdr bytes are compared with the reference's own stbi_write_hdr in tests (oracle/_ref):
// header text, bottom-up row flip, RGBE mantissas by truncation, per-channel run-length coding
// with literal blocks <= 128 and runs of 3..127 — all identical.
// The PNG uses stored (uncompressed) deflate blocks so no zlib is needed; any decoder reads it.
#include <cmath>
#include <cstdio>
#include <cstring>

#include "host.hpp"

namespace mcpt::host {
namespace {

void to_rgbe(const float* c, unsigned char out[4]) {
	float m = c[0] > (c[1] > c[2] ? c[1] : c[2]) ? c[0] : (c[1] > c[2] ? c[1] : c[2]);
	if (m < 1e-32f) { out[0] = out[1] = out[2] = out[3] = 0; return; }
	int e;
	float scale = float(std::frexp(m, &e)) * 256.0f / m;
	out[0] = (unsigned char)(c[0] * scale);
	out[1] = (unsigned char)(c[1] * scale);
	out[2] = (unsigned char)(c[2] * scale);
	out[3] = (unsigned char)(e + 128);
}

// One channel plane of a scanline -> RLE bytes.
void rle_plane(const unsigned char* p, int w, std::vector<unsigned char>& out) {
	int x = 0;
	while (x < w) {
		// locate the next position where three equal bytes start
		int run = x;
		while (run + 2 < w && !(p[run] == p[run + 1] && p[run] == p[run + 2])) ++run;
		const bool have_run = run + 2 < w;
		const int lit_end = have_run ? run : w;
		while (x < lit_end) {                 // literal block(s)
			int len = lit_end - x;
			if (len > 128) len = 128;
			out.push_back((unsigned char)len);
			out.insert(out.end(), p + x, p + x + len);
			x += len;
		}
		if (have_run) {
			int end = run;
			while (end < w && p[end] == p[x]) ++end;
			while (x < end) {
				int len = end - x;
				if (len > 127) len = 127;
				out.push_back((unsigned char)(len + 128));
				out.push_back(p[x]);
				x += len;
			}
		}
	}
}

uint32_t crc_table[256];
bool crc_ready = false;
uint32_t crc32(uint32_t crc, const unsigned char* d, size_t n) {
	if (!crc_ready) {
		for (uint32_t i = 0; i < 256; ++i) {
			uint32_t c = i;
			for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
			crc_table[i] = c;
		}
		crc_ready = true;
	}
	crc = ~crc;
	for (size_t i = 0; i < n; ++i) crc = crc_table[(crc ^ d[i]) & 0xFF] ^ (crc >> 8);
	return ~crc;
}
void be32(std::vector<unsigned char>& v, uint32_t x) {
	v.push_back((unsigned char)(x >> 24)); v.push_back((unsigned char)(x >> 16));
	v.push_back((unsigned char)(x >> 8)); v.push_back((unsigned char)x);
}
void png_chunk(std::vector<unsigned char>& f, const char tag[4], const std::vector<unsigned char>& body) {
	be32(f, uint32_t(body.size()));
	size_t start = f.size();
	f.insert(f.end(), tag, tag + 4);
	f.insert(f.end(), body.begin(), body.end());
	be32(f, crc32(0, f.data() + start, f.size() - start));
}

}  // namespace

bool write_hdr(const std::string& path, int w, int h, const float* rgba) {
	if (w <= 0 || h <= 0 || !rgba) return false;
	FILE* f = std::fopen(path.c_str(), "wb");
	if (!f) return false;
	std::fputs("#?RADIANCE\n# Written by stb_image_write.h\nFORMAT=32-bit_rle_rgbe\n", f);
	std::fprintf(f, "EXPOSURE=          1.0000000000000\n\n-Y %d +X %d\n", h, w);
	std::vector<unsigned char> planes(size_t(w) * 4), line;
	for (int row = 0; row < h; ++row) {
		const float* src = rgba + size_t(4) * size_t(w) * size_t(h - 1 - row);   // bottom-up flip
		line.clear();
		if (w < 8 || w >= 32768) {
			for (int x = 0; x < w; ++x) {
				unsigned char px[4];
				to_rgbe(src + 4 * x, px);
				line.insert(line.end(), px, px + 4);
			}
		} else {
			for (int x = 0; x < w; ++x) {
				unsigned char px[4];
				to_rgbe(src + 4 * x, px);
				for (int c = 0; c < 4; ++c) planes[size_t(c) * size_t(w) + size_t(x)] = px[c];
			}
			line.push_back(2); line.push_back(2);
			line.push_back((unsigned char)((w & 0xff00) >> 8));
			line.push_back((unsigned char)(w & 0xff));
			for (int c = 0; c < 4; ++c) rle_plane(planes.data() + size_t(c) * size_t(w), w, line);
		}
		if (std::fwrite(line.data(), 1, line.size(), f) != line.size()) { std::fclose(f); return false; }
	}
	return std::fclose(f) == 0;
}

bool write_png(const std::string& path, int w, int h, const float* rgba) {
	if (w <= 0 || h <= 0 || !rgba) return false;
	// raw scanlines: filter byte 0 + RGB8, top row first (rgba row 0 is the image bottom)
	std::vector<unsigned char> raw;
	raw.reserve(size_t(h) * (size_t(w) * 3 + 1));
	for (int row = 0; row < h; ++row) {
		const float* src = rgba + size_t(4) * size_t(w) * size_t(h - 1 - row);
		raw.push_back(0);
		for (int x = 0; x < w; ++x)
			for (int c = 0; c < 3; ++c) {
				float v = src[4 * x + c];
				v = v > 0.0f ? std::pow(v, 1 / 2.2f) : 0.0f;     // display gamma, testkernel.cl:11-13
				if (!(v < 1.0f)) v = 1.0f;
				raw.push_back((unsigned char)(v * 255.0f + 0.5f));
			}
	}
	std::vector<unsigned char> z;
	z.push_back(0x78); z.push_back(0x01);
	uint32_t a = 1, b = 0;
	size_t pos = 0;
	while (pos < raw.size() || raw.empty()) {
		size_t len = raw.size() - pos;
		if (len > 65535) len = 65535;
		bool final_block = pos + len == raw.size();
		z.push_back(final_block ? 1 : 0);
		z.push_back((unsigned char)(len & 0xFF)); z.push_back((unsigned char)(len >> 8));
		z.push_back((unsigned char)(~len & 0xFF)); z.push_back((unsigned char)((~len >> 8) & 0xFF));
		z.insert(z.end(), raw.begin() + long(pos), raw.begin() + long(pos + len));
		for (size_t i = pos; i < pos + len; ++i) { a = (a + raw[i]) % 65521u; b = (b + a) % 65521u; }
		pos += len;
		if (final_block) break;
	}
	be32(z, (b << 16) | a);
	std::vector<unsigned char> file = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
	std::vector<unsigned char> ihdr;
	be32(ihdr, uint32_t(w)); be32(ihdr, uint32_t(h));
	ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
	png_chunk(file, "IHDR", ihdr);
	png_chunk(file, "IDAT", z);
	png_chunk(file, "IEND", {});
	FILE* f = std::fopen(path.c_str(), "wb");
	if (!f) return false;
	bool ok = std::fwrite(file.data(), 1, file.size(), f) == file.size();
	return (std::fclose(f) == 0) && ok;
}

}  // namespace mcpt::host
