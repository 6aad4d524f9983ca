// gather_kernel.cu — the per-bin read set.
//
// Reference code replaced: loadFilteredReads, src/d_mapper.h:393-443.  After the filter has produced
// origReadIdMap[bin] (m read ids, or pair ids for paired-end), the reference copies, read by read,
//     [ seqs[id]            for the m ids ]                      forward reads (first mates)
//     [ seqs[id + numPairs] for the m ids ]                      second mates            (paired-end only)
//     [ seqs[x + numReads]  for every x pushed so far ]          reverse complements of all of the above
// into the child mapper's read set and appends the shifted ids to origReadIdMap[bin] in that order.  The main
// read set holds the reverse complements explicitly (appendReverseComplement, store_seqs.h:271-290); here only
// the forward half is resident and the reverse complements are produced on the fly, byte-exact for Dna5Q
// (complement keeps the quality, N stays 252: modifier_functors.h:276-284, alphabet_residue.h:357-365).
#include "dyibf_internal.h"

namespace dyibf {

__device__ __forceinline__ void gather_source(const GatherArgs& a, uint32_t j, uint32_t& fwd_index, bool& rc, uint32_t& id_out)
{
    const uint32_t sec = j / a.m, i = j - sec * a.m;
    const uint32_t id = a.ids[i];
    if (a.group == 4) {                       // paired-end: n_fwd = 2p
        const uint32_t p = a.n_fwd / 2;
        fwd_index = id + ((sec & 1u) ? p : 0u);
        rc = sec >= 2;
        id_out = id + sec * p;                // id, id+p, id+2p, id+3p  (numPairs = p, numReads = 2p)
    } else {                                  // single-end: n_fwd = n
        fwd_index = id;
        rc = sec >= 1;
        id_out = id + sec * a.n_fwd;          // id, id+numReads
    }
}

__global__ void __launch_bounds__(256) gather_plan_kernel(GatherArgs a, uint64_t* __restrict__ lens, uint32_t* __restrict__ out_ids)
{
    const uint64_t n_out = (uint64_t)a.m * a.group;
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_out; j += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t fi, id; bool rc;
        gather_source(a, (uint32_t)j, fi, rc, id);
        lens[j] = a.offsets[fi + 1] - a.offsets[fi];
        out_ids[j] = id;
    }
}

__device__ __forceinline__ uint8_t complement_byte(uint8_t b, uint32_t enc)
{
    if (enc == DYIBF_ENC_DNA5Q) return b >= 252 ? (uint8_t)252 : (uint8_t)((b & 0xFCu) | (3u - (b & 3u)));
    if (enc == DYIBF_ENC_DNA5)  return b >= 4 ? (uint8_t)4 : (uint8_t)(3u - b);
    switch (b) {
        case 'A': return 'T'; case 'C': return 'G'; case 'G': return 'C'; case 'T': case 'U': return 'A';
        case 'a': return 't'; case 'c': return 'g'; case 'g': return 'c'; case 't': case 'u': return 'a';
        default:  return 'N';
    }
}

// one warp per output sequence
__global__ void __launch_bounds__(256) gather_copy_kernel(GatherArgs a, const uint64_t* __restrict__ limits, uint8_t* __restrict__ out)
{
    const uint64_t n_out = (uint64_t)a.m * a.group;
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t j = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; j < n_out; j += n_warps) {
        uint32_t fi, id; bool rc;
        gather_source(a, (uint32_t)j, fi, rc, id);
        const uint64_t o0 = a.offsets[fi], L = a.offsets[fi + 1] - o0;
        const uint8_t* __restrict__ src = a.bytes + o0;
        uint8_t* __restrict__ dst = out + limits[j];
        if (!rc) for (uint64_t x = lane; x < L; x += 32) dst[x] = src[x];
        else     for (uint64_t x = lane; x < L; x += 32) dst[x] = complement_byte(src[L - 1 - x], a.encoding);
    }
}

cudaError_t launch_gather_plan(const GatherArgs& a, uint64_t* lens, uint32_t* out_ids, cudaStream_t st)
{
    const uint64_t n_out = (uint64_t)a.m * a.group;
    if (n_out == 0) return cudaSuccess;
    uint64_t blocks = (n_out + 255) / 256;
    if (blocks > sm_count_current() * 8) blocks = sm_count_current() * 8;
    gather_plan_kernel<<<(unsigned)blocks, 256, 0, st>>>(a, lens, out_ids);
    return cudaGetLastError();
}

cudaError_t launch_gather_copy(const GatherArgs& a, const uint64_t* limits, uint8_t* out_bytes, cudaStream_t st)
{
    const uint64_t n_out = (uint64_t)a.m * a.group;
    if (n_out == 0) return cudaSuccess;
    uint64_t blocks = (n_out + 7) / 8;
    if (blocks > sm_count_current() * 16) blocks = sm_count_current() * 16;
    gather_copy_kernel<<<(unsigned)blocks, 256, 0, st>>>(a, limits, out_bytes);
    return cudaGetLastError();
}

} // namespace dyibf
