// ibf_hash.cuh — device-side k-mer value and hash-position arithmetic, bit-identical to the reference.
//
//   k-mer value : SeqAn Shape<Dna,SimpleShape> rolling hash (include/seqan/index/shape_base.h:479-494,
//                 699-714) in closed form: H_i = sum_j d_{i+j} 4^(k-1-j) + [c_0 == N] 4^(k+i)  (mod 2^64),
//                 d = code & 3 (N hashes as A).  With the 2-bit big-endian packing of SeqMeta this is a
//                 funnel shift of two neighbouring words — no rolling state, any lane can compute any k-mer.
//   positions   : d_bloom_filter.h:183-184,286-291: v = pre[i]*H; v ^= v >> 27; v %= P; block v.
#pragma once
#include "dyibf_internal.h"

namespace dyibf {

__device__ __forceinline__ uint64_t kmer_value(const uint64_t* __restrict__ w, uint32_t i, uint32_t k, uint32_t firstN)
{
    const uint32_t q = i >> 5, s = (i & 31u) * 2u;
    const uint64_t hi = w[q], lo = w[q + 1];
    uint64_t v = s ? ((hi << s) | (lo >> (64u - s))) : hi;
    v >>= (64u - 2u * k);
    // leading-N term: the raw ordValue 4 of a leading N is never subtracted again and rides up by 4x per
    // k-mer until it leaves the 64-bit word (shape_base.h:489 vs :492)
    if (firstN && (k + i) < 32u) v += 1ull << (2u * (k + i));
    return v;
}

// same from two already-loaded words
__device__ __forceinline__ uint64_t kmer_value2(uint64_t hi, uint64_t lo, uint32_t i, uint32_t k, uint32_t firstN)
{
    const uint32_t s = (i & 31u) * 2u;
    uint64_t v = s ? ((hi << s) | (lo >> (64u - s))) : hi;
    v >>= (64u - 2u * k);
    if (firstN && (k + i) < 32u) v += 1ull << (2u * (k + i));
    return v;
}

// v mod P, exact for every u64 v.  Power-of-two P: mask.  Otherwise q' = mulhi(v, floor(2^64/P)) is q or q-1.
__device__ __forceinline__ uint64_t mod_blocks(uint64_t v, const HashParams& hp)
{
    if (hp.pow2) return v & (hp.P - 1);
    uint64_t q = __umul64hi(v, hp.magic);
    uint64_t r = v - q * hp.P;
    while (r >= hp.P) r -= hp.P;
    return r;
}

__device__ __forceinline__ uint64_t block_index(uint64_t kmer, uint32_t i, const HashParams& hp)
{
    if (hp.direct) return kmer < hp.P ? kmer : hp.sink;             // d_kdx_filter.h:171-175,287-291
    uint64_t v = hp.pre[i] * kmer;
    v ^= v >> 27;
    return mod_blocks(v, hp);
}

} // namespace dyibf
