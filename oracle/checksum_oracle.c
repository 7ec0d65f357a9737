/*
 * CPU oracle for the probe kernels -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * The probes are NEW capability (the reference launches no GPU work), so there is no reference code to
 * restate: this file is the plain-C statement of what include/gfd_b200.h promises for
 * gfd_hbm_fill_pattern / gfd_hbm_read_checksum / gfd_hbm_copy, written independently of the CUDA code
 * (byte-at-a-time word assembly, no vectorisation tricks) so the two can be compared bit for bit.
 * Parity for these is therefore "pinned by definition", not by reference vectors -- see DESIGN.md.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.
 */
#include <stddef.h>
#include <stdint.h>

#define GOLDEN 0x9E3779B97F4A7C15ull

/* word[i] = GOLDEN * (i + seed), little-endian bytes; a ragged tail keeps the low bytes of the last word */
void oracle_fill_pattern(uint8_t *buf, size_t bytes, uint64_t seed) {
    for (size_t b = 0; b < bytes; b++) {
        uint64_t w = GOLDEN * ((uint64_t)(b >> 3) + seed);
        buf[b] = (uint8_t)(w >> (8 * (b & 7)));
    }
}

/* xor64 / sum64 over little-endian 64-bit words, tail zero-padded to 8 bytes */
void oracle_checksum(const uint8_t *buf, size_t bytes, uint64_t out[2]) {
    uint64_t x = 0, s = 0;
    size_t b = 0;
    while (b < bytes) {
        uint64_t w = 0;
        for (int k = 0; k < 8 && b < bytes; k++, b++) w |= (uint64_t)buf[b] << (8 * k);
        x ^= w;
        s += w;
    }
    out[0] = x;
    out[1] = s;
}

/* closed form of the pattern's checksum without materialising the buffer (whole words only) */
void oracle_pattern_checksum(size_t words, uint64_t seed, uint64_t out[2]) {
    uint64_t x = 0, s = 0;
    for (size_t i = 0; i < words; i++) {
        uint64_t w = GOLDEN * ((uint64_t)i + seed);
        x ^= w;
        s += w;
    }
    out[0] = x;
    out[1] = s;
}
