// Host helpers for the frame-list ingest (the reference reads its stage-1 checkpoint with
// bz2.BZ2File + pickle, Utils/utilities.py:84-96).  ingest.py decodes the bzip2 blocks of the file in parallel;
// what stayed serial there was finding the blocks (the 48-bit block magic at ANY bit offset: sixteen
// bytes.find passes over the file under the GIL, 3.4 s for a 337 MB file) and re-aligning each block's
// bits with numpy.  Both are plain byte loops here: one pass with a 64-bit shift register, split over
// threads, and a bit-granular copy.  No CUDA in this file.
#include <algorithm>
#include <stdint.h>
#include <thread>
#include <vector>
#include "../../include/rti_b200.h"

namespace {

// start bytes [a, b): every bit offset 8 j + s (s = 0..7, MSB first) at which the 48 bits equal `magic`.
// The pattern shifted by s has bytes j + 1 and j + 2 of its window fully determined for every s, so two
// 256-entry tables say for which shifts a position can match at all (about one position in 8 000); only
// those are compared in full.
void scan_range(const uint8_t *buf, int64_t n, int64_t a, int64_t b, uint64_t magic, std::vector<int64_t> &hits) {
    const uint64_t mask48 = (1ull << 48) - 1;
    uint8_t cand1[256] = {}, cand2[256] = {};
    for (int s = 0; s < 8; ++s) {
        cand1[(uint8_t)((magic << (8 - s)) >> 40)] |= (uint8_t)(1u << s);        // window byte 1
        cand2[(uint8_t)((magic << (8 - s)) >> 32)] |= (uint8_t)(1u << s);        // window byte 2
    }
    auto check = [&](int64_t j, unsigned m) {
        uint64_t win = 0;                                        // bytes j .. j + 6, zeros beyond the buffer
        for (int k = 0; k < 7; ++k) win = (win << 8) | (j + k < n ? buf[j + k] : 0);
        for (int s = 0; s < 8; ++s)
            if ((m >> s & 1) && ((win >> (8 - s)) & mask48) == magic && j * 8 + s + 48 <= n * 8)
                hits.push_back(j * 8 + s);
    };
    const int64_t end = std::min(std::min(n, b), n - 2);         // positions whose bytes j + 1, j + 2 exist
    for (int64_t j = a; j < end; ++j) {
        const unsigned m = cand1[buf[j + 1]] & cand2[buf[j + 2]];
        if (m) check(j, m);
    }
    // (a pattern needs at least 6 bytes, so start bytes n - 2 and n - 1 cannot hold one)
}

}  // namespace

// Bit offsets (ascending) of every occurrence of the 48-bit value `magic48` in buf[0, n), at any bit alignment.
// Returns how many there are; only the first `cap` are written.  threads <= 0: one per core, at most 32.
extern "C" int64_t rti_find_magic48(const uint8_t *buf, int64_t n, uint64_t magic48, int64_t *bit_offsets,
                                    int64_t cap, int threads) {
    if (!buf || n < 6 || (!bit_offsets && cap > 0)) return 0;
    if (threads <= 0) threads = (int)std::min(32u, std::max(1u, std::thread::hardware_concurrency()));
    const int64_t min_chunk = 1 << 20;
    int parts = (int)std::min<int64_t>(threads, std::max<int64_t>(1, n / min_chunk));
    std::vector<std::vector<int64_t>> found((size_t)parts);
    std::vector<std::thread> th;
    for (int k = 0; k < parts; ++k) {
        const int64_t a = n * k / parts, b = n * (k + 1) / parts;
        if (parts == 1) scan_range(buf, n, a, b, magic48, found[0]);
        else th.emplace_back([=, &found] { scan_range(buf, n, a, b, magic48, found[(size_t)k]); });
    }
    for (auto &t : th) t.join();
    int64_t count = 0;
    for (auto &v : found)
        for (int64_t off : v) {
            if (count < cap) bit_offsets[count] = off;
            ++count;
        }
    return count;
}

// Copies nbits bits (MSB first within a byte, bzip2's order) from bit offset src_bit0 of src to bit offset
// dst_bit0 of dst.  Bits of dst's first byte before dst_bit0 are kept, bits of its last byte after the end are
// zeroed.  Returns the bit offset just past what was written.
extern "C" int64_t rti_copy_bits(const uint8_t *src, int64_t src_bit0, int64_t nbits, uint8_t *dst, int64_t dst_bit0) {
    if (!src || !dst || nbits <= 0) return dst_bit0;
    uint64_t acc = 0;                      // pending output bits, the oldest highest
    int accbits = 0;
    int64_t di = dst_bit0 >> 3, si = src_bit0 >> 3, remaining = nbits;
    const int dlead = (int)(dst_bit0 & 7), slead = (int)(src_bit0 & 7);
    if (dlead) {
        acc = dst[di] >> (8 - dlead);
        accbits = dlead;
    }
    auto flush = [&] {
        while (accbits >= 8) {
            dst[di++] = (uint8_t)(acc >> (accbits - 8));
            accbits -= 8;
        }
    };
    if (slead) {
        const int take = (int)std::min<int64_t>(8 - slead, remaining);
        const uint64_t v = (uint64_t)((src[si] & (0xFFu >> slead)) >> (8 - slead - take));
        acc = (acc << take) | v;
        accbits += take;
        remaining -= take;
        ++si;
        flush();
    }
    while (remaining >= 8) {
        acc = ((acc << 8) | src[si++]) & 0xFFFF;
        accbits += 8;
        remaining -= 8;
        flush();
    }
    if (remaining) {
        acc = (acc << remaining) | (uint64_t)(src[si] >> (8 - remaining));
        accbits += (int)remaining;
        flush();
    }
    if (accbits) dst[di] = (uint8_t)(acc << (8 - accbits));
    return dst_bit0 + nbits;
}
