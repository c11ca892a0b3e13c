// 128-bit content hash for cache keys and integrity checks (scene cache, derived-layout blobs).  Not cryptographic.
// Data is cut into fixed 1 MiB blocks hashed independently (two multiply-mix lanes over 8-byte words) and the block
// digests are folded in order, so the result does not depend on how many threads computed the blocks.
#pragma once
#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace mcpt {

struct Hash128 {
	uint64_t a = 0x9E3779B97F4A7C15ull, b = 0xC2B2AE3D27D4EB4Full;
	bool operator==(const Hash128& o) const { return a == o.a && b == o.b; }
	bool operator!=(const Hash128& o) const { return !(*this == o); }
};

namespace hash_detail {
inline uint64_t mix(uint64_t x) {
	x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull;
	x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull;
	return x ^ (x >> 32);
}
inline Hash128 block(const unsigned char* p, size_t n) {
	uint64_t a = 0x243F6A8885A308D3ull ^ n, b = 0x13198A2E03707344ull + n;
	size_t i = 0;
	for (; i + 8 <= n; i += 8) {
		uint64_t w;
		std::memcpy(&w, p + i, 8);
		a = (a ^ w) * 0x9FB21C651E98DF25ull; a = (a << 29) | (a >> 35);
		b = (b + w) * 0xA0761D6478BD642Full; b ^= b >> 31;
	}
	uint64_t w = 0;
	if (i < n) std::memcpy(&w, p + i, n - i);
	a = (a ^ w) * 0x9FB21C651E98DF25ull;
	b = (b + w) * 0xA0761D6478BD642Full;
	return Hash128{mix(a ^ (b >> 7)), mix(b ^ (a << 9))};
}
}  // namespace hash_detail

// Folds `n` bytes into `h` (order matters: h = fold(h, data)).
inline void hash_update(Hash128& h, const void* data, size_t n, unsigned max_threads = 8) {
	constexpr size_t BLOCK = size_t(1) << 20;
	const unsigned char* p = static_cast<const unsigned char*>(data);
	const size_t nblocks = (n + BLOCK - 1) / BLOCK;
	std::vector<Hash128> d(nblocks);
	auto run = [&](size_t lo, size_t hi) {
		for (size_t k = lo; k < hi; ++k) d[k] = hash_detail::block(p + k * BLOCK, std::min(BLOCK, n - k * BLOCK));
	};
	const unsigned nt = unsigned(std::min<size_t>(std::max(1u, std::min(max_threads, std::thread::hardware_concurrency())), nblocks / 4 + 1));
	if (nt <= 1) run(0, nblocks);
	else {
		std::vector<std::thread> th;
		for (unsigned t = 0; t < nt; ++t) th.emplace_back(run, nblocks * t / nt, nblocks * (t + 1) / nt);
		for (auto& x : th) x.join();
	}
	h.a = hash_detail::mix(h.a ^ (uint64_t(n) * 0x9E3779B97F4A7C15ull));
	h.b = hash_detail::mix(h.b + uint64_t(n));
	for (const Hash128& x : d) {
		h.a = hash_detail::mix(h.a ^ x.a) + x.b;
		h.b = hash_detail::mix(h.b + x.b) ^ x.a;
	}
}
inline void hash_update(Hash128& h, const char* s) { hash_update(h, s, std::strlen(s) + 1, 1); }

}  // namespace mcpt
