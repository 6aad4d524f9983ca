/*
 * bench_support/synth.c — deterministic synthetic binned genomes and reads for bench.py and
 * the full-size parity tests.  Host code, OpenMP; not part of the product library and not part of the oracle.
 *
 * Genomes: per bin, uniform ACGT from splitmix64 (seed 0xD2EA3A2A00000000 + bin, 32 bases per draw), one run of
 * N (length 1..50) per 10 000 bases (rate 1e-4 runs/base), and bins with bin % 64 == 7 start with N (exercises
 * the leading-N term of the reference hash).
 * Reads: seed 0xBEEF0000 + cfg.  90 %: uniform (bin, position, strand), copied, 1 %/base substitutions, one 1-bp
 * indel in 5 % of the reads; 10 %: uniform random (unmappable).  0.5 % of all reads get 1..3 N, with positions 0
 * and L-1 among the choices.
 * All sequences are Dna5 codes (A,C,G,T,N = 0..4), one byte per base.
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <immintrin.h>

static inline uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

#define GENOME_SEED 0xD2EA3A2A00000000ULL

/* bases [0, len) of bin `bin` */
void synth_genome(uint32_t bin, uint64_t len, uint8_t *out)
{
    const uint64_t seed = splitmix64(GENOME_SEED + bin);
    #pragma omp parallel for schedule(static)
    for (uint64_t blk = 0; blk < (len + 31) / 32; ++blk) {
        uint64_t r = splitmix64(seed ^ (blk * 0x2545F4914F6CDD1DULL));
        uint64_t end = (blk + 1) * 32 < len ? (blk + 1) * 32 : len;
        if (end - blk * 32 == 32) {                /* 8 bases per store: bits 2j..2j+1 -> byte j */
            for (int q = 0; q < 4; ++q) {
                uint64_t v = _pdep_u64(r >> (16 * q), 0x0303030303030303ULL);
                memcpy(out + blk * 32 + 8 * q, &v, 8);
            }
            continue;
        }
        for (uint64_t i = blk * 32; i < end; ++i) { out[i] = (uint8_t)(r & 3); r >>= 2; }
    }
    /* N runs: one per 10 000-base block */
    const uint64_t nblk = len / 10000;
    #pragma omp parallel for schedule(static)
    for (uint64_t j = 0; j < nblk; ++j) {
        uint64_t r = splitmix64(seed ^ 0xA5A5A5A5ULL ^ (j * 0x9E3779B97F4A7C15ULL));
        uint64_t start = j * 10000 + r % 9950, n = 1 + (r >> 32) % 50;
        for (uint64_t i = start; i < start + n && i < len; ++i) out[i] = 4;
    }
    if (len && bin % 64 == 7) out[0] = 4;
}

/* all bins, each `len` bases, concatenated: out[bin * len + i] */
void synth_genomes(uint32_t bins, uint64_t len, uint8_t *out)
{
    for (uint32_t b = 0; b < bins; ++b) synth_genome(b, len, out + (uint64_t)b * len);
}

/* reads [first, first+n) of configuration `cfg`, each L bases, from genomes laid out as synth_genomes does.
 * out[(r-first)*L + i].  origin (optional): the bin each read was drawn from, 0xffffffff for random reads. */
void synth_reads(const uint8_t *genomes, uint32_t bins, uint64_t glen, uint32_t cfg, uint64_t first, uint64_t n,
                 uint32_t L, uint8_t *out, uint32_t *origin)
{
    const uint64_t seed = splitmix64(0xBEEF0000ULL + cfg);
    #pragma omp parallel for schedule(static)
    for (uint64_t q = 0; q < n; ++q) {
        const uint64_t r = first + q;
        uint64_t s = splitmix64(seed ^ (r * 0xD6E8FEB86659FD93ULL));
        uint8_t *o = out + q * L;
        uint8_t buf[1024 + 2];
        uint32_t src_bin = 0xffffffffu;
        if (L > 1024 || glen < (uint64_t)L + 2) { memset(o, 0, L); if (origin) origin[q] = src_bin; continue; }
        if (s % 10 != 0) {
            uint64_t a = splitmix64(s + 1), b = splitmix64(s + 2);
            src_bin = (uint32_t)(a % bins);
            uint64_t pos = b % (glen - L - 1);
            int rc = (int)((a >> 40) & 1);
            const uint8_t *g = genomes + (uint64_t)src_bin * glen + pos;
            uint32_t take = L;
            int indel = 0; uint32_t ipos = 0;
            if ((b >> 40) % 20 == 0) { indel = ((b >> 50) & 1) ? 1 : -1; ipos = 1 + (uint32_t)((b >> 20) % (L - 2)); }
            if (indel < 0) {            /* deletion: read L+1 reference bases, drop one */
                take = L + 1;
                for (uint32_t i = 0, j = 0; i < take; ++i) if (i != ipos) buf[j++] = g[i];
            } else if (indel > 0) {     /* insertion: read L-1 reference bases, insert one */
                take = L - 1;
                for (uint32_t i = 0, j = 0; i < take; ++i) { if (i == ipos) buf[j++] = (uint8_t)((b >> 60) & 3); buf[j++] = g[i]; }
            } else memcpy(buf, g, L);
            /* substitutions, 1 % per base */
            uint64_t t = splitmix64(s + 3);
            for (uint32_t i = 0; i < L; ++i) {
                t = t * 6364136223846793005ULL + 1442695040888963407ULL;
                if ((t >> 33) % 100 == 0) buf[i] = (uint8_t)((t >> 20) & 3);
            }
            if (rc) for (uint32_t i = 0; i < L; ++i) { uint8_t c = buf[L - 1 - i]; o[i] = c >= 4 ? 4 : (uint8_t)(3 - c); }
            else memcpy(o, buf, L);
        } else {
            uint64_t t = splitmix64(s + 4);
            for (uint32_t i = 0; i < L; ++i) {
                if ((i & 31) == 0) t = splitmix64(t + i);
                o[i] = (uint8_t)((t >> (2 * (i & 31))) & 3);
            }
        }
        uint64_t nn = splitmix64(s + 5);
        if (nn % 200 == 0) {            /* 0.5 %: 1..3 N; choices include position 0 and L-1 */
            uint32_t cnt = 1 + (uint32_t)((nn >> 8) % 3);
            for (uint32_t j = 0; j < cnt; ++j) {
                uint64_t w = splitmix64(nn + j);
                uint32_t pos = (w % 4 == 0) ? 0 : (w % 4 == 1) ? L - 1 : (uint32_t)((w >> 8) % L);
                o[pos] = 4;
            }
        }
        if (origin) origin[q] = src_bin;
    }
}

/* paired-end reads (cfg3): pairs [first, first+n), mates of L bases each.  90 %: a fragment of length
 * L + 50 .. 2L + 250 at a uniform (bin, position); mate 1 = its first L bases, mate 2 = the reverse complement of
 * its last L bases (FR orientation), the whole fragment flipped for half of the pairs; 1 %/base substitutions in
 * both mates.  10 %: two uniform random mates.  0.5 % of the mates get 1..3 N.  m1[(p-first)*L + i], m2 likewise. */
void synth_pairs(const uint8_t *genomes, uint32_t bins, uint64_t glen, uint32_t cfg, uint64_t first, uint64_t n,
                 uint32_t L, uint8_t *m1, uint8_t *m2)
{
    const uint64_t seed = splitmix64(0xBEEF0000ULL + cfg);
    #pragma omp parallel for schedule(static)
    for (uint64_t q = 0; q < n; ++q) {
        const uint64_t r = first + q;
        const uint64_t s = splitmix64(seed ^ (r * 0xD6E8FEB86659FD93ULL));
        uint8_t *o[2] = { m1 + q * L, m2 + q * L };
        if (L > 1024 || glen < 3ull * L + 300) { memset(o[0], 0, L); memset(o[1], 0, L); continue; }
        if (s % 10 != 0) {
            const uint64_t a = splitmix64(s + 1), b = splitmix64(s + 2);
            const uint32_t bin = (uint32_t)(a % bins);
            const uint32_t frag = L + 50 + (uint32_t)((a >> 32) % (L + 201));
            const uint64_t pos = b % (glen - frag - 1);
            const uint8_t *g = genomes + (uint64_t)bin * glen + pos;
            const int flip = (int)((a >> 61) & 1);
            /* left end forward, right end reverse-complemented; flip swaps which mate is which */
            uint8_t *left = o[flip], *right = o[1 - flip];
            memcpy(left, g, L);
            for (uint32_t i = 0; i < L; ++i) { uint8_t c = g[frag - 1 - i]; right[i] = c >= 4 ? 4 : (uint8_t)(3 - c); }
            uint64_t t = splitmix64(s + 3);
            for (int m = 0; m < 2; ++m)
                for (uint32_t i = 0; i < L; ++i) {
                    t = t * 6364136223846793005ULL + 1442695040888963407ULL;
                    if ((t >> 33) % 100 == 0) o[m][i] = (uint8_t)((t >> 20) & 3);
                }
        } else {
            uint64_t t = splitmix64(s + 4);
            for (int m = 0; m < 2; ++m)
                for (uint32_t i = 0; i < L; ++i) {
                    if ((i & 31) == 0) t = splitmix64(t + i + 1000 * m);
                    o[m][i] = (uint8_t)((t >> (2 * (i & 31))) & 3);
                }
        }
        for (int m = 0; m < 2; ++m) {
            uint64_t nn = splitmix64(s + 5 + m);
            if (nn % 200 == 0) {
                uint32_t cnt = 1 + (uint32_t)((nn >> 8) % 3);
                for (uint32_t j = 0; j < cnt; ++j) {
                    uint64_t w = splitmix64(nn + j);
                    uint32_t pos = (w % 4 == 0) ? 0 : (w % 4 == 1) ? L - 1 : (uint32_t)((w >> 8) % L);
                    o[m][pos] = 4;
                }
            }
        }
    }
}
