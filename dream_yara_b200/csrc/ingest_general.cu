// ingest_general.cu — FASTQ in SeqAn's full generality on the device: sequences and qualities wrapped over several lines,
// blank lines between records, '@' and '+' inside quality strings.  The fall-back of dyibf_parse_records for texts the
// four-line tokeniser of ingest_kernel.cu refuses.
//
// Rules restated (seqan seq_io/fasta_fastq.h:302-345, stream/tokenization.h:426-477; oracle/ibf_oracle.c:381-417 is the CPU
// restatement pinned against the compiled reference): a record is '@' + the rest of that line (id); every non-white-space
// character up to the next '+' (anywhere) is sequence and must be in the alphabet; the rest of the '+' line is skipped; then
// as many non-white-space characters as the sequence has are qualities (the end of the text may cut them short); white space
// is skipped and the next character must be '@' (or the text ends).  Text before the first '@' is skipped.
//
// Where a record ends depends on its sequence length, so record starts are not a prefix function of the text.  They are
// found speculatively instead:
//   1 candidates : every '@' that follows white space (and the first '@' of the text) may open a record;     2 reads
//   2 walk       : one thread per candidate parses "its" record as if it were one: where the next record would start,
//                  how many sequence characters, the first error;                                             ~1 read
//   3 chain      : next-pointers -> candidate indices (binary search), then pointer jumping from the first candidate marks
//                  the candidates that really are records (log2(candidates) rounds over two small arrays)
//   4 scan       : record numbers and sequence offsets of the real records
//   5 emit       : one thread per real record copies its sequence characters and writes its limit.
// A real record whose successor starts right behind its qualities without white space in between is not a candidate: such a
// text (SeqAn accepts it) is reported as E_FQ_NOSPACE instead of being mis-parsed.  Throughput is that of byte-wise walks
// (tens of GB/s), an order of magnitude below the four-line path: this is the path for unusual files, not the fast one.
#include "dyibf_handle.h"

namespace dyibf {

namespace {

const int GT = 256;                        // threads per block
const int GB = 16;                         // bytes per thread in the candidate passes
const uint64_t GTILE = (uint64_t)GT * GB;
const uint32_t FULLM = 0xffffffffu;

enum { E_ALPHABET = 1, E_FQ_LAYOUT = 2, E_FQ_QUAL = 3, E_FQ_TRUNC = 4, E_FQ_NOSPACE = 5 };

struct FqgRec { uint64_t next; uint64_t seq_begin; uint32_t nseq; uint32_t err; };

__device__ __forceinline__ bool g_ws(uint8_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
__device__ __forceinline__ bool g_nl(uint8_t c) { return c == '\n' || c == '\r'; }
__device__ __forceinline__ bool g_alpha(uint8_t c, int iupac)
{
    const uint8_t u = c & 0xdfu;           // lower case -> upper case; nothing else becomes a letter
    if (u == 'A' || u == 'C' || u == 'G' || u == 'T' || u == 'N') return true;
    if (!iupac) return false;
    return u == 'U' || u == 'W' || u == 'Y' || u == 'M' || u == 'H' || u == 'K' || u == 'R' || u == 'D' || u == 'S' || u == 'B' || u == 'V';
}
__device__ __forceinline__ bool g_candidate(const uint8_t* __restrict__ t, uint64_t p, uint64_t first_at)
{
    return t[p] == '@' && (p == first_at || (p > 0 && g_ws(t[p - 1])));
}

__global__ void fqg_init(unsigned long long* g) { g[0] = ~0ull; g[1] = ~0ull; }     // first '@', smallest error word

__global__ void __launch_bounds__(GT) fqg_first_at(const uint8_t* __restrict__ t, uint64_t n, unsigned long long* __restrict__ g)
{
    const uint64_t p0 = ((uint64_t)blockIdx.x * GT + threadIdx.x) * GB;
    unsigned long long best = ~0ull;
    for (int j = GB - 1; j >= 0; --j) if (p0 + j < n && t[p0 + j] == '@') best = p0 + j;
    #pragma unroll
    for (int d = 16; d; d >>= 1) { const unsigned long long o = __shfl_xor_sync(FULLM, best, d); best = o < best ? o : best; }
    if ((threadIdx.x & 31) == 0 && best != ~0ull) atomicMin(g, best);
}

// candidates per tile (WRITE = false) / their positions in text order (WRITE = true; tile_off = exclusive scan of the counts)
template <bool WRITE>
__global__ void __launch_bounds__(GT) fqg_candidates(const uint8_t* __restrict__ t, uint64_t n, const unsigned long long* __restrict__ g,
                                                     uint64_t* __restrict__ tile_cnt, uint64_t* __restrict__ cand_pos)
{
    __shared__ uint32_t warp_sum[GT / 32];
    const uint64_t first_at = g[0];
    const uint64_t p0 = ((uint64_t)blockIdx.x * GT + threadIdx.x) * GB;
    uint32_t mask = 0;
    #pragma unroll
    for (int j = 0; j < GB; ++j) if (p0 + j < n && g_candidate(t, p0 + j, first_at)) mask |= 1u << j;
    const uint32_t mine = __popc(mask), lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = mine;
    #pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(FULLM, incl, d); if (lane >= (uint32_t)d) incl += o; }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    uint32_t before = 0, total = 0;
    #pragma unroll
    for (int w = 0; w < GT / 32; ++w) { if (w < (int)warp) before += warp_sum[w]; total += warp_sum[w]; }
    if (!WRITE) { if (threadIdx.x == 0) tile_cnt[blockIdx.x] = total; return; }
    uint64_t k = tile_cnt[blockIdx.x] + before + incl - mine;
    for (uint32_t m = mask; m; m &= m - 1) cand_pos[k++] = p0 + (uint32_t)(__ffs(m) - 1);
}

// one thread per candidate: the record that would start there
__global__ void __launch_bounds__(GT) fqg_walk(const uint8_t* __restrict__ t, uint64_t n, int iupac, const uint64_t* __restrict__ cand_pos,
                                               uint64_t n_cand, FqgRec* __restrict__ rec)
{
    const uint64_t c = (uint64_t)blockIdx.x * GT + threadIdx.x;
    if (c >= n_cand) return;
    uint64_t i = cand_pos[c] + 1;                                           // skipOne('@')
    while (i < n && !g_nl(t[i])) ++i;                                       // readLine(meta)
    if (i < n) { if (t[i] == '\r' && i + 1 < n && t[i + 1] == '\n') ++i; ++i; }
    FqgRec r; r.seq_begin = i; r.nseq = 0; r.err = 0; r.next = n;
    uint64_t len = 0;
    for (; i < n && t[i] != '+'; ++i) {                                     // readUntil(seq, '+', IgnoreOrAssert)
        const uint8_t ch = t[i];
        if (g_ws(ch)) continue;
        if (!g_alpha(ch, iupac)) { r.err = E_ALPHABET; r.next = i; break; }
        ++len;
    }
    if (!r.err) {
        if (i >= n) { r.err = E_FQ_TRUNC; r.next = n; }                     // skipOne('+') at the end: UnexpectedEnd
        else if (len > 0xffffffffull) { r.err = E_FQ_LAYOUT; r.next = i; }
        else {
            ++i;
            while (i < n && !g_nl(t[i])) ++i;                               // skipLine
            if (i < n) { if (t[i] == '\r' && i + 1 < n && t[i + 1] == '\n') ++i; ++i; }
            uint64_t left = len;                                            // CountDownFunctor<Not<IsWhitespace>>(length(seq))
            for (; i < n && left; ++i) if (!g_ws(t[i])) --left;
            while (i < n && g_ws(t[i])) ++i;                                // skipUntil(Not<IsWhitespace>)
            if (i < n && t[i] != '@') { r.err = E_FQ_QUAL; r.next = i; }    // "quality string is expected to be of the same length"
            else r.next = i;
            r.nseq = (uint32_t)len;
        }
    }
    rec[c] = r;
}

// next position -> candidate index (n_cand = the end of the chain)
__global__ void __launch_bounds__(GT) fqg_link(const uint64_t* __restrict__ cand_pos, uint64_t n_cand, uint64_t n, FqgRec* __restrict__ rec,
                                               uint32_t* __restrict__ J)
{
    const uint64_t c = (uint64_t)blockIdx.x * GT + threadIdx.x;
    if (c >= n_cand) return;
    uint32_t j = (uint32_t)n_cand;
    if (!rec[c].err && rec[c].next < n) {
        const uint64_t want = rec[c].next;
        uint64_t lo = c + 1, hi = n_cand;                                   // the next record lies behind this one
        while (lo < hi) { const uint64_t mid = (lo + hi) >> 1; if (cand_pos[mid] < want) lo = mid + 1; else hi = mid; }
        if (lo < n_cand && cand_pos[lo] == want) j = (uint32_t)lo;
        else rec[c].err = E_FQ_NOSPACE;                                     // an '@' right behind the qualities: not a candidate
    }
    J[c] = j;
}

__global__ void __launch_bounds__(GT) fqg_mark(const uint32_t* __restrict__ J, uint8_t* __restrict__ mark, uint64_t n_cand)
{
    const uint64_t c = (uint64_t)blockIdx.x * GT + threadIdx.x;
    if (c >= n_cand || !mark[c]) return;
    const uint32_t j = J[c];
    if (j < n_cand) mark[j] = 1;
}

__global__ void __launch_bounds__(GT) fqg_jump(const uint32_t* __restrict__ J, uint32_t* __restrict__ J2, uint64_t n_cand)
{
    const uint64_t c = (uint64_t)blockIdx.x * GT + threadIdx.x;
    if (c >= n_cand) return;
    const uint32_t j = J[c];
    J2[c] = j < n_cand ? J[j] : (uint32_t)n_cand;
}

__global__ void __launch_bounds__(GT) fqg_flags(const uint8_t* __restrict__ mark, const FqgRec* __restrict__ rec, uint64_t n_cand,
                                                uint64_t* __restrict__ rec_flag, uint64_t* __restrict__ seq_len, unsigned long long* __restrict__ g)
{
    const uint64_t c = (uint64_t)blockIdx.x * GT + threadIdx.x;
    if (c >= n_cand) return;
    const bool real = mark[c] != 0;
    const FqgRec r = rec[c];
    // a record in error ends the chain and is not a record of the output (SeqAn throws there)
    rec_flag[c] = real && !r.err ? 1 : 0;
    seq_len[c] = real && !r.err ? r.nseq : 0;
    if (real && r.err) atomicMin(&g[1], ((unsigned long long)r.next << 8) | r.err);
}

__global__ void __launch_bounds__(GT) fqg_emit(const uint8_t* __restrict__ t, uint64_t n, const uint8_t* __restrict__ mark, const FqgRec* __restrict__ rec,
                                               uint64_t n_cand, const uint64_t* __restrict__ rec_idx, const uint64_t* __restrict__ seq_off,
                                               uint8_t* __restrict__ out_seqs, uint64_t* __restrict__ out_limits, uint64_t limit_base)
{
    const uint64_t c = (uint64_t)blockIdx.x * GT + threadIdx.x;
    if (c == 0) out_limits[rec_idx[n_cand]] = limit_base + seq_off[n_cand];                  // the totals close the limits
    if (c >= n_cand || !mark[c] || rec[c].err) return;
    uint64_t o = seq_off[c];
    out_limits[rec_idx[c]] = limit_base + o;
    for (uint64_t i = rec[c].seq_begin; i < n && t[i] != '+'; ++i) if (!g_ws(t[i])) out_seqs[o++] = t[i];
}

inline unsigned blocks_for(uint64_t n) { return (unsigned)((n + GT - 1) / GT); }

} // namespace

// Passes 1-4 (everything but the output).  *err_word: (position << 8 | code) of the first error in text order, ~0 if none.
// The object keeps its device arrays for emit().
int FastqGeneral::count(const uint8_t* text, uint64_t n, int iupac, cudaStream_t st, uint64_t* n_records, uint64_t* n_bases, uint64_t* err_word)
{
    text_ = text; n_ = n; n_cand_ = 0;
    *n_records = 0; *n_bases = 0; *err_word = ~0ull;
    if (n == 0) return DYIBF_OK;
    const uint64_t n_tiles = (n + GTILE - 1) / GTILE;
    CU(globals_.reserve(64));
    CU(tiles_.reserve(scan_workspace_elems(n_tiles) * 8));
    unsigned long long* g = static_cast<unsigned long long*>(globals_.p);
    uint64_t* tile_cnt = static_cast<uint64_t*>(tiles_.p);
    fqg_init<<<1, 1, 0, st>>>(g);
    fqg_first_at<<<(unsigned)n_tiles, GT, 0, st>>>(text, n, g);
    fqg_candidates<false><<<(unsigned)n_tiles, GT, 0, st>>>(text, n, g, tile_cnt, nullptr);
    CU(cudaGetLastError());
    CU(launch_scan_u64(tile_cnt, n_tiles, st));
    uint64_t n_cand = 0;
    CU(cudaMemcpyAsync(&n_cand, tile_cnt + n_tiles, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (n_cand == 0) { *err_word = (n << 8) | E_FQ_LAYOUT; return DYIBF_OK; }       // no '@' at all: the caller reports UnexpectedEnd
    if (n_cand >= 0xffffffffull) return fail(DYIBF_ERR_ARG, "dyibf_parse_records: too many FASTQ record candidates in one text");
    n_cand_ = n_cand;
    CU(cand_.reserve(n_cand * 8));
    CU(rec_.reserve(n_cand * sizeof(FqgRec)));
    CU(jump_.reserve(2 * n_cand * 4));
    CU(mark_.reserve(n_cand));
    CU(scan_a_.reserve(scan_workspace_elems(n_cand) * 8));
    CU(scan_b_.reserve(scan_workspace_elems(n_cand) * 8));
    uint64_t* cand_pos = static_cast<uint64_t*>(cand_.p);
    FqgRec* rec = static_cast<FqgRec*>(rec_.p);
    uint32_t* J = static_cast<uint32_t*>(jump_.p); uint32_t* J2 = J + n_cand;
    uint8_t* mark = static_cast<uint8_t*>(mark_.p);
    fqg_candidates<true><<<(unsigned)n_tiles, GT, 0, st>>>(text, n, g, tile_cnt, cand_pos);
    fqg_walk<<<blocks_for(n_cand), GT, 0, st>>>(text, n, iupac, cand_pos, n_cand, rec);
    fqg_link<<<blocks_for(n_cand), GT, 0, st>>>(cand_pos, n_cand, n, rec, J);
    CU(cudaMemsetAsync(mark, 0, n_cand, st));
    CU(cudaMemsetAsync(mark, 1, 1, st));                                     // the first candidate is the first '@' of the text
    for (uint64_t span = 1; ; span <<= 1) {                                  // after the round with J = next^span: everything within 2 span - 1
        fqg_mark<<<blocks_for(n_cand), GT, 0, st>>>(J, mark, n_cand);
        if (span >= n_cand) break;
        fqg_jump<<<blocks_for(n_cand), GT, 0, st>>>(J, J2, n_cand);
        uint32_t* tmp = J; J = J2; J2 = tmp;
    }
    uint64_t* rec_flag = static_cast<uint64_t*>(scan_a_.p); uint64_t* seq_len = static_cast<uint64_t*>(scan_b_.p);
    fqg_flags<<<blocks_for(n_cand), GT, 0, st>>>(mark, rec, n_cand, rec_flag, seq_len, g);
    CU(cudaGetLastError());
    CU(launch_scan_u64(rec_flag, n_cand, st));
    CU(launch_scan_u64(seq_len, n_cand, st));
    unsigned long long gg[2] = {0, 0};
    CU(cudaMemcpyAsync(n_records, rec_flag + n_cand, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(n_bases, seq_len + n_cand, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(gg, g, 16, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    *err_word = gg[1];
    return DYIBF_OK;
}

// Pass 5: out_seqs[n_bases], out_limits[n_records + 1] (device pointers)
int FastqGeneral::emit(uint8_t* out_seqs, uint64_t* out_limits, uint64_t limit_base, cudaStream_t st)
{
    if (n_cand_ == 0) {
        CU(cudaMemcpyAsync(out_limits, &limit_base, 8, cudaMemcpyHostToDevice, st));
        return DYIBF_OK;
    }
    fqg_emit<<<blocks_for(n_cand_), GT, 0, st>>>(text_, n_, static_cast<const uint8_t*>(mark_.p), static_cast<const FqgRec*>(rec_.p), n_cand_,
                                                 static_cast<const uint64_t*>(scan_a_.p), static_cast<const uint64_t*>(scan_b_.p),
                                                 out_seqs, out_limits, limit_base);
    CU(cudaGetLastError());
    return DYIBF_OK;
}

FastqGeneral::~FastqGeneral()
{
    globals_.release(); tiles_.release(); cand_.release(); rec_.release(); jump_.release(); mark_.release(); scan_a_.release(); scan_b_.release();
}

} // namespace dyibf
