// copy_pool_test.cpp — TEST ONLY: exercises the product's host copy pool (csrc/copy_pool.h) on the
// CPU: every size class (inline, split, ragged), concurrent callers, and content equality.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "copy_pool.h"

static bool one(size_t bytes, unsigned seed) {
    std::vector<unsigned char> src(bytes + 64), dst(bytes + 64, 0xEE);
    for (size_t i = 0; i < src.size(); ++i) src[i] = (unsigned char)((i * 2654435761u + seed) >> 13);
    b200v::CopyPool::instance().copy(dst.data() + 32, src.data() + 32, bytes);
    if (bytes && std::memcmp(dst.data() + 32, src.data() + 32, bytes) != 0) return false;
    for (int i = 0; i < 32; ++i)                       // nothing outside the range was written
        if (dst[i] != 0xEE || dst[32 + bytes + i] != 0xEE) return false;
    return true;
}

int main() {
    const size_t sizes[] = {0, 1, 4095, 1 << 20, (2 << 20) - 1, 2 << 20, (2 << 20) + 1, (16 << 20) + 12345, 64 << 20};
    for (size_t s : sizes)
        if (!one(s, (unsigned)s)) { std::printf("FAIL size %zu\n", s); return 1; }
    // concurrent callers (the per-device host threads of a fan-out): calls serialise inside the pool
    bool ok[4] = {false, false, false, false};
    std::vector<std::thread> th;
    for (int t = 0; t < 4; ++t) th.emplace_back([&, t] { ok[t] = one((8 << 20) + 7 * t, 100 + t) && one(12345 + t, t); });
    for (auto &t : th) t.join();
    for (bool b : ok) if (!b) { std::printf("FAIL concurrent\n"); return 1; }
    std::printf("COPY_POOL_OK threads=%d\n", b200v::CopyPool::instance().threads());
    return 0;
}
