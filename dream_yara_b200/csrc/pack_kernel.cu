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

struct SeqLoc { uint64_t pos; uint64_t len; const uint8_t* bytes; };

__device__ __forceinline__ uint64_t seq_pos(const PackSource& S, uint64_t s, uint64_t n_in)
{
    if (s >= n_in) {                               // one past the end
        uint32_t g = S.n_ranges - 1;
        return S.pack_base[g] + S.offsets[g][S.cnt] - S.first_off[g];
    }
    uint32_t g = (uint32_t)(s / S.cnt), i = (uint32_t)(s - (uint64_t)g * S.cnt);
    return S.pack_base[g] + S.offsets[g][i] - S.first_off[g];
}
__device__ __forceinline__ uint64_t word_off_of(const PackSource& S, uint64_t s, uint64_t n_in)
{
    return seq_pos(S, s, n_in) / 32 + 2 * s;
}

template <int ENC>
__global__ void __launch_bounds__(256) pack_kernel(PackSource S, int derive_rc, uint64_t* __restrict__ packed,
                                                   SeqMeta* __restrict__ meta, uint64_t rc_word_base, uint64_t total_words)
{
    const uint64_t n_in = (uint64_t)S.n_ranges * S.cnt;
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t tile = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; tile * 32 < total_words; tile += n_warps) {
        const uint64_t t0 = tile * 32, t = t0 + lane;
        // largest s with word_off(s) <= t0 (word_off(0) == 0)
        uint64_t lo = 0, hi = n_in;                // invariant: word_off(lo) <= t0 < word_off(hi) (hi may be n_in)
        while (hi - lo > 1) {
            uint64_t mid = (lo + hi) >> 1;
            if (word_off_of(S, mid, n_in) <= t0) lo = mid; else hi = mid;
        }
        uint64_t s = lo;
        if (t >= total_words || n_in == 0) continue;
        while (s + 1 < n_in && word_off_of(S, s + 1, n_in) <= t) ++s;

        const uint32_t g = (uint32_t)(s / S.cnt), i = (uint32_t)(s - (uint64_t)g * S.cnt);
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
            #pragma unroll 4
            for (uint32_t x = 0; x < nb; ++x) {
                uint32_t c = decode_base(src[b0 + x], ENC);
                fw |= (uint64_t)(c & 3u) << (62 - 2 * x);                 // N (4) -> 00
            }
            if (derive_rc) {
                #pragma unroll 4
                for (uint32_t x = 0; x < nb; ++x) {
                    uint32_t c = decode_base(src[L - 1 - (b0 + x)], ENC);
                    uint32_t r = c >= 4 ? 0u : 3u - c;                    // complement; N stays N -> 00
                    rw |= (uint64_t)r << (62 - 2 * x);
                }
            }
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
    if (blocks > 148ull * 64) blocks = 148ull * 64;
    dim3 grid((unsigned)blocks), block(256);
    switch (encoding) {
        case DYIBF_ENC_DNA5:  pack_kernel<DYIBF_ENC_DNA5><<<grid, block, 0, st>>>(src, derive_rc, packed, meta, rc_word_base, total_words); break;
        case DYIBF_ENC_DNA5Q: pack_kernel<DYIBF_ENC_DNA5Q><<<grid, block, 0, st>>>(src, derive_rc, packed, meta, rc_word_base, total_words); break;
        default:              pack_kernel<DYIBF_ENC_ASCII><<<grid, block, 0, st>>>(src, derive_rc, packed, meta, rc_word_base, total_words); break;
    }
    return cudaGetLastError();
}

} // namespace dyibf
