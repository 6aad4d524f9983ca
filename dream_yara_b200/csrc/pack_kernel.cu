// pack_kernel.cu — sequence bytes -> 2-bit packed words + per-sequence meta, forward and (optionally)
// reverse-complement strands, on the device.
//
// Replaces, for the filter's purposes, the host-side read set the reference's mapper builds with
// loadReads + appendReverseComplement (src/mapper.h:444-469, src/store_seqs.h:271-290) and the
// Iupac->Dna5 conversion addFastaFile performs (src/d_bloom_filter.h:127-147,
// seqan basic/alphabet_residue_tabs.h:270-290).  The complement of N is N (modifier_functors.h:170) and N
// hashes as A (alphabet_residue.h:680-683), so both strands store N as 00 and only "first base is N" survives
// as a flag (it feeds the leading-N term of the k-mer value, see ibf_hash.cuh).
//
// Work decomposition: one thread per packed output word; a warp owns 32 consecutive words of the packed
// buffer, finds the sequence its first word belongs to with one binary search over the (monotone) word
// offsets and every lane walks forward from there.  That keeps 150 bp reads and 100 Mbp contigs on the same
// code path with full lane utilisation.
#include "dyibf_internal.h"

namespace dyibf {

__device__ __forceinline__ uint32_t decode_base(uint8_t b, int enc)
{
    if (enc == DYIBF_ENC_DNA5)  return b > 4 ? 4u : (uint32_t)b;
    if (enc == DYIBF_ENC_DNA5Q) return b >= 252 ? 4u : (uint32_t)(b & 3);
    switch (b) {                                   // ASCII through Iupac -> Dna5
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': case 'U': case 'u': return 3;
        default: return 4;
    }
}

// ---- 32 bases at a time ----------------------------------------------------------------------------------------------
// 0x80 in every byte of x that equals ch (exact, no borrow between bytes)
__device__ __forceinline__ uint32_t eq4(uint32_t x, uint32_t ch)
{
    const uint32_t d = x ^ (ch * 0x01010101u);
    return ~(((d & 0x7f7f7f7fu) + 0x7f7f7f7fu) | d) & 0x80808080u;
}
// 0x80 in every non-zero byte of x
__device__ __forceinline__ uint32_t nonzero4(uint32_t x) { return (((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u; }

// four sequence bytes -> four 2-bit codes (one per byte; N and everything that decodes to N -> 0), as decode_base does
template <int ENC>
__device__ __forceinline__ uint32_t codes4(uint32_t x)
{
    uint32_t n_flags, code;                                            // n_flags: 0x80 where the byte is N
    if (ENC == DYIBF_ENC_DNA5)       { n_flags = nonzero4(x & 0xfcfcfcfcu); code = x & 0x03030303u; }            // > 3: N (4 and above)
    else if (ENC == DYIBF_ENC_DNA5Q) { n_flags = 0x80808080u ^ nonzero4(~x & 0xfcfcfcfcu); code = x & 0x03030303u; }   // >= 252: N
    else {
        const uint32_t u = x & 0xdfdfdfdfu;                            // upper case
        n_flags = 0x80808080u ^ (eq4(u, 'A') | eq4(u, 'C') | eq4(u, 'G') | eq4(u, 'T') | eq4(u, 'U'));
        const uint32_t v = (x >> 1) & 0x03030303u;                     // A 0, C 1, T/U 2, G 3
        code = v ^ ((v >> 1) & 0x01010101u);                           // A 0, C 1, G 2, T/U 3
    }
    return code & ~((n_flags >> 7) * 0xffu);
}

// nb (1..32) sequence bytes at p (any alignment) -> 2 bits per base, first base in the two top bits; bases >= nb are 00.
// Reads the aligned 32-bit words that hold the bytes: up to 3 bytes on either side, inside the same words.
template <int ENC>
__device__ __forceinline__ uint64_t pack32(const uint8_t* __restrict__ p, uint32_t nb, bool complement)
{
    const uint32_t a = (uint32_t)(reinterpret_cast<uintptr_t>(p) & 3u);
    const uint32_t* __restrict__ q = reinterpret_cast<const uint32_t*>(p - a);
    const uint32_t n_words = (a + nb + 3u) >> 2;                        // <= 9
    uint32_t t[9];
    #pragma unroll
    for (int i = 0; i < 9; ++i) t[i] = (uint32_t)i < n_words ? __ldg(q + i) : 0u;
    uint64_t out = 0;
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uint32_t x = __funnelshift_r(t[i], t[i + 1], 8 * a);      // bytes 4i .. 4i+3 of the sequence window
        uint32_t c = codes4<ENC>(x);
        if (complement) {                                               // 3 - c; N stays N (00)
            uint32_t n_keep;
            if (ENC == DYIBF_ENC_DNA5)       n_keep = nonzero4(x & 0xfcfcfcfcu);
            else if (ENC == DYIBF_ENC_DNA5Q) n_keep = 0x80808080u ^ nonzero4(~x & 0xfcfcfcfcu);
            else { const uint32_t u = x & 0xdfdfdfdfu; n_keep = 0x80808080u ^ (eq4(u, 'A') | eq4(u, 'C') | eq4(u, 'G') | eq4(u, 'T') | eq4(u, 'U')); }
            c = (c ^ 0x03030303u) & ~((n_keep >> 7) * 0xffu);
        }
        // bytes past nb carry whatever follows the sequence: cleared
        const int left = (int)nb - 4 * i;
        const uint32_t keep = left >= 4 ? 0xffffffffu : (left <= 0 ? 0u : (0xffffffffu >> (8 * (4 - left))));
        c &= keep;
        // byte k of c (base 4i+k) -> bits 7-2k .. 6-2k of one byte: multiply gathers the four 2-bit fields into the top byte
        const uint32_t byte = (c * 0x40100401u) >> 24;
        out |= (uint64_t)byte << (56 - 8 * i);
    }
    return out;
}

// reverse the order of the 32 2-bit fields of x
__device__ __forceinline__ uint64_t reverse_bases(uint64_t x)
{
    const uint64_t r = __brevll(x);                                     // fields reversed, bits inside a field swapped
    return ((r & 0x5555555555555555ull) << 1) | ((r >> 1) & 0x5555555555555555ull);
}

struct SeqLoc { uint64_t pos; uint64_t len; const uint8_t* bytes; };

template <bool ONE>
__device__ __forceinline__ uint64_t seq_pos(const PackSource& S, uint64_t s)
{
    if (ONE) return S.pack_base[0] + S.offsets[0][s] - S.first_off[0];        // s == cnt: one past the end
    uint32_t g = (uint32_t)s / S.cnt;                                                       // n_in < 2^32: 32-bit division
    if (g >= S.n_ranges) g = S.n_ranges - 1;                                                // one past the end
    const uint32_t i = (uint32_t)s - g * S.cnt;
    return S.pack_base[g] + S.offsets[g][i] - S.first_off[g];
}
template <bool ONE>
__device__ __forceinline__ uint64_t word_off_of(const PackSource& S, uint64_t s)
{
    return seq_pos<ONE>(S, s) / 32 + 2 * s;
}

// ONE: the batch is a single byte range (n_ranges == 1): no division, no dynamically indexed kernel parameters
template <int ENC, bool ONE>
__global__ void __launch_bounds__(256) pack_kernel(PackSource S, int derive_rc, uint64_t* __restrict__ packed,
                                                   SeqMeta* __restrict__ meta, uint64_t rc_word_base, uint64_t total_words)
{
    const uint64_t n_in = (uint64_t)S.n_ranges * S.cnt;
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    const double per_word = total_words ? (double)n_in / (double)total_words : 0.0;
    for (uint64_t tile = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; tile * 32 < total_words; tile += n_warps) {
        const uint64_t t0 = tile * 32, t = t0 + lane;
        // largest s with word_off(s) <= t0 (word_off(0) == 0).  Reads of one length make word_off linear in s, so the
        // proportional guess is right or next to it: gallop from there (two loads instead of a 23-step dependent binary
        // search over 10 M reads), then bisect what is left.
        uint64_t lo = 0, hi = n_in;                // invariant: word_off(lo) <= t0 < word_off(hi) (hi may be n_in)
        if (n_in > 1) {
            uint64_t guess = (uint64_t)((double)t0 * per_word);
            if (guess >= n_in) guess = n_in - 1;
            if (word_off_of<ONE>(S, guess) <= t0) {
                lo = guess;
                for (uint64_t step = 1; ; step *= 2) {
                    const uint64_t nx = lo + step < n_in ? lo + step : n_in;
                    if (nx == n_in || word_off_of<ONE>(S, nx) > t0) { hi = nx; break; }
                    lo = nx;
                }
            } else {
                hi = guess;
                for (uint64_t step = 1; ; step *= 2) {
                    const uint64_t nx = hi > step ? hi - step : 0;
                    if (nx == 0 || word_off_of<ONE>(S, nx) <= t0) { lo = nx; break; }
                    hi = nx;
                }
            }
        }
        while (hi - lo > 1) {
            uint64_t mid = (lo + hi) >> 1;
            if (word_off_of<ONE>(S, mid) <= t0) lo = mid; else hi = mid;
        }
        uint64_t s = lo;
        if (t >= total_words || n_in == 0) continue;
        while (s + 1 < n_in && word_off_of<ONE>(S, s + 1) <= t) ++s;

        const uint32_t g = ONE ? 0u : (uint32_t)s / S.cnt, i = (uint32_t)s - g * S.cnt;
        const uint64_t o0 = S.offsets[g][i], o1 = S.offsets[g][i + 1];
        const uint64_t L = o1 - o0;
        const uint8_t* __restrict__ src = S.bytes[g] + o0;
        const uint64_t woff = (S.pack_base[g] + o0 - S.first_off[g]) / 32 + 2 * s;
        const uint64_t j = t - woff;               // word of this sequence
        const uint64_t nw = (L + 31) / 32;

        uint64_t fw = 0, rw = 0;
        if (j < nw) {
            const uint64_t b0 = j * 32;
            const uint32_t nb = (uint32_t)((L - b0) < 32 ? (L - b0) : 32);
            fw = pack32<ENC>(src + b0, nb, false);
            if (derive_rc)                                             // bases L-1-b0 down to L-b0-nb, complemented
                rw = reverse_bases(pack32<ENC>(src + (L - b0 - nb), nb, true)) << (2 * (32 - nb));
        }
        packed[t] = fw;                             // also zeroes the pad word and any gap words
        if (derive_rc) packed[rc_word_base + t] = rw;
        if (j == 0) {
            SeqMeta m;
            m.word_off = woff; m.len = (uint32_t)L;
            m.flags = (L > 0 && decode_base(src[0], ENC) == 4) ? 1u : 0u;
            meta[s] = m;
            if (derive_rc) {
                m.word_off = rc_word_base + woff;
                m.flags = (L > 0 && decode_base(src[L - 1], ENC) == 4) ? 1u : 0u;
                meta[n_in + s] = m;
            }
        }
    }
}

cudaError_t launch_pack(const PackSource& src, int encoding, int derive_rc, uint64_t* packed, SeqMeta* meta,
                        uint64_t total_words, uint64_t rc_word_base, cudaStream_t st)
{
    // total_words = packed_words_for(total bytes, n_in): the forward strand's share of `packed`; the
    // reverse-complement strand (derive_rc) occupies the same shape starting at rc_word_base.
    if (total_words == 0) return cudaSuccess;
    uint64_t tiles = (total_words + 31) / 32;
    uint64_t blocks = (tiles + 7) / 8;
    if (blocks > sm_count_current() * 64) blocks = sm_count_current() * 64;
    dim3 grid((unsigned)blocks), block(256);
#define DYIBF_PACK(E) do { \
        if (src.n_ranges == 1) pack_kernel<E, true><<<grid, block, 0, st>>>(src, derive_rc, packed, meta, rc_word_base, total_words); \
        else                   pack_kernel<E, false><<<grid, block, 0, st>>>(src, derive_rc, packed, meta, rc_word_base, total_words); } while (0)
    switch (encoding) {
        case DYIBF_ENC_DNA5:  DYIBF_PACK(DYIBF_ENC_DNA5); break;
        case DYIBF_ENC_DNA5Q: DYIBF_PACK(DYIBF_ENC_DNA5Q); break;
        default:              DYIBF_PACK(DYIBF_ENC_ASCII); break;
    }
#undef DYIBF_PACK
    return cudaGetLastError();
}

} // namespace dyibf
