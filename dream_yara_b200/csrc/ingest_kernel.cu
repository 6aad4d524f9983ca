// ingest_kernel.cu — FASTA / FASTQ text -> sequence characters + limits on the device (dyibf_parse_records).
//
// Replaces, for the step in front of the filter, SeqAn's sequential tokenizer as the reference drives it:
// readRecords(me.names, me.seqs, fileIn) in loadReads (src/mapper.h:451, src/store_seqs.h:175-188) and the readRecord
// loop of addFastaFile (src/d_bloom_filter.h:127-147); rules from seqan seq_io/fasta_fastq.h:268-282 (Fasta),
// :302-345 (Fastq), stream/tokenization.h:51-58 (IsInAlphabet), :145-149 (whitespace = ' ' '\t' '\n' '\r').
//
// Both formats are state machines whose state at a byte is a prefix function of the text, so the text is cut into tiles
// of 4 KiB (256 threads x 16 bytes) and parsed in five streaming passes:
//   1 summary : per tile, what a later tile needs to know about it (below);       one read of the text
//   2 scan    : tile summaries -> state at the start of every tile                (n/4096 elements)
//   3 count   : per tile, records started / sequence characters / quality characters, and validation;  second read
//   4 scan    : exclusive prefix sums of those counts over the tiles
//   5 emit    : sequence characters to out_seqs, limits[r] at every record start;  third read
// FASTA  — SeqAn: skip to the next '>', the rest of that line is the id, then every non-whitespace character up to the
//   next '>' (anywhere, not only at a line start) is sequence.  An id ends at a newline and a '>' inside an id is text, so
//   "inside an id" at byte i <=> the last of {'\n', '\r', '>'} before i is a '>': the tile summary is the kind of its last
//   such character, the scan propagates the last non-empty summary.  Characters before the first '>' are skipped.
// FASTQ  — the four-line layout ('@' id / sequence / '+' / qualities, one line each): the state is the number of '\n'
//   before the byte, mod 4; the tile summary is its newline count, the scan a prefix sum.  SeqAn also accepts sequences
//   and qualities wrapped over several lines and blank lines between records; those need the sequential reader (the
//   number of quality characters to consume is the length of the sequence read so far) and are REJECTED here with
//   DYIBF_ERR_FORMAT instead of being mis-parsed: every '@' line must open a line whose number is 0 mod 4, every third
//   line must open with '+', and at every end of a quality line the running totals of sequence and quality characters
//   must agree — which is also SeqAn's own check (:341-344).  Trailing whitespace after the last record is ignored.
// Characters outside the target alphabet are SeqAn ParseErrors (reads: ACGTN, contigs: the 16 IUPAC letters; either
// case); here the smallest offending offset is reported through dyibf_last_error.
#include "dyibf_internal.h"

#include <cstdio>

namespace dyibf {

namespace {

const uint32_t FULL = 0xffffffffu;
const int      IT = 256;                   // threads per tile
const int      IB = 16;                    // bytes per thread
const uint64_t TILE = (uint64_t)IT * IB;

enum { G_BEGIN = 0, G_END = 1, G_ERR = 2, G_N = 4 };     // device globals: first '>' / effective end / first error
enum { E_ALPHABET = 1, E_FQ_LAYOUT = 2, E_FQ_QUAL = 3, E_FQ_TRUNC = 4 };
enum { D_NONE = 0, D_NL = 1, D_GT = 2 };

__device__ __forceinline__ bool is_ws(uint32_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
__device__ __forceinline__ bool is_nl(uint32_t c) { return c == '\n' || c == '\r'; }

// IsInAlphabet<Dna5Q> / IsInAlphabet<Iupac>: the character, upper-cased, is one of the alphabet's letters
__device__ __forceinline__ bool in_alphabet(uint32_t c, int iupac)
{
    const uint32_t u = (c >= 'a' && c <= 'z') ? c - 32 : c;
    if (u < 'A' || u > 'Z') return false;
    const uint32_t dna5 = (1u << ('A' - 'A')) | (1u << ('C' - 'A')) | (1u << ('G' - 'A')) | (1u << ('T' - 'A')) | (1u << ('N' - 'A'));
    const uint32_t more = (1u << ('U' - 'A')) | (1u << ('W' - 'A')) | (1u << ('Y' - 'A')) | (1u << ('M' - 'A')) | (1u << ('H' - 'A')) |
                          (1u << ('K' - 'A')) | (1u << ('R' - 'A')) | (1u << ('D' - 'A')) | (1u << ('S' - 'A')) | (1u << ('B' - 'A')) |
                          (1u << ('V' - 'A'));
    return ((iupac ? (dna5 | more) : dna5) >> (u - 'A')) & 1u;
}

// the 16 bytes of one thread, in registers
struct Chunk {
    uint32_t w[4];
    uint32_t m;                // valid bytes (0..16)
    __device__ __forceinline__ uint32_t at(int j) const { return (w[j >> 2] >> (8 * (j & 3))) & 0xffu; }
};

__device__ __forceinline__ Chunk load_chunk(const uint8_t* __restrict__ text, uint64_t i0, uint64_t end, bool aligned)
{
    Chunk c; c.w[0] = c.w[1] = c.w[2] = c.w[3] = 0;
    c.m = i0 >= end ? 0u : (uint32_t)(end - i0 < (uint64_t)IB ? end - i0 : (uint64_t)IB);
    if (c.m == IB && aligned) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(text + i0));
        c.w[0] = v.x; c.w[1] = v.y; c.w[2] = v.z; c.w[3] = v.w;
    } else {
        for (uint32_t j = 0; j < c.m; ++j) c.w[j >> 2] |= (uint32_t)__ldg(text + i0 + j) << (8 * (j & 3));
    }
    return c;
}

struct OpSum  { __device__ __forceinline__ uint64_t operator()(uint64_t a, uint64_t b) const { return a + b; } };
struct OpLast { __device__ __forceinline__ uint64_t operator()(uint64_t a, uint64_t b) const { return b ? b : a; } };   // a earlier

// exclusive scan over the 256 threads of a block (identity 0); *total = combination of all
template <typename Op>
__device__ __forceinline__ uint64_t block_excl_scan(uint64_t v, Op op, uint64_t* sm /* [IT/32] */, uint64_t* total)
{
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint64_t x = v;
    #pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint64_t y = __shfl_up_sync(FULL, x, d);
        if (lane >= (uint32_t)d) x = op(y, x);
    }
    __syncthreads();                                   // sm may still be read by a previous scan
    if (lane == 31) sm[warp] = x;
    __syncthreads();
    if (warp == 0) {
        uint64_t w = lane < IT / 32 ? sm[lane] : 0;
        #pragma unroll
        for (int d = 1; d < IT / 32; d <<= 1) {
            const uint64_t y = __shfl_up_sync(FULL, w, d);
            if (lane >= (uint32_t)d) w = op(y, w);
        }
        if (lane < IT / 32) sm[lane] = w;
    }
    __syncthreads();
    uint64_t excl = __shfl_up_sync(FULL, x, 1);
    if (lane == 0) excl = 0;
    const uint64_t prefix = warp ? sm[warp - 1] : 0;
    *total = sm[IT / 32 - 1];
    return op(prefix, excl);
}

// ---- pass 1 ----------------------------------------------------------------------------------------------------------
template <int FMT>
__global__ void __launch_bounds__(IT) ingest_summary_kernel(const uint8_t* __restrict__ text, uint64_t n, bool aligned,
                                                            uint64_t* __restrict__ summary, unsigned long long* __restrict__ g)
{
    __shared__ uint64_t sm[IT / 32];
    const uint64_t i0 = (uint64_t)blockIdx.x * TILE + (uint64_t)threadIdx.x * IB;
    const Chunk c = load_chunk(text, i0, n, aligned);
    uint64_t v = 0, total;
    if (FMT == DYIBF_FMT_FASTA) {
        uint32_t last = D_NONE, first_gt = IB;
        #pragma unroll
        for (int j = 0; j < IB; ++j) {
            if ((uint32_t)j >= c.m) break;
            const uint32_t ch = c.at(j);
            if (ch == '>') { last = D_GT; if (first_gt == IB) first_gt = j; }
            else if (is_nl(ch)) last = D_NL;
        }
        // the block's first '>' is the one of the lowest thread that saw any
        const uint32_t any = __ballot_sync(FULL, first_gt != IB);
        if (any && (threadIdx.x & 31) == (uint32_t)(__ffs(any) - 1)) atomicMin(&g[G_BEGIN], (unsigned long long)(i0 + first_gt));
        block_excl_scan(last, OpLast(), sm, &total);
        v = total;
    } else {
        uint32_t nl = 0, last_graph = 0;
        #pragma unroll
        for (int j = 0; j < IB; ++j) {
            if ((uint32_t)j >= c.m) break;
            const uint32_t ch = c.at(j);
            nl += ch == '\n';
            if (!is_ws(ch)) last_graph = j + 1;
        }
        const uint64_t e = last_graph ? i0 + last_graph : 0;            // one past this thread's last non-whitespace byte
        uint64_t emax = e;
        #pragma unroll
        for (int d = 16; d; d >>= 1) { const uint64_t o = __shfl_xor_sync(FULL, emax, d); emax = o > emax ? o : emax; }
        if ((threadIdx.x & 31) == 0 && emax) atomicMax(&g[G_END], (unsigned long long)emax);
        block_excl_scan(nl, OpSum(), sm, &total);
        v = total;
    }
    if (threadIdx.x == 0) summary[blockIdx.x] = v;
}

// ---- pass 2 (FASTA): entry[t] = kind of the last delimiter in tiles [0, t), single block ------------------------------
__global__ void __launch_bounds__(1024) ingest_propagate_kernel(uint64_t* __restrict__ a, uint64_t n)
{
    __shared__ uint64_t part[1024];
    const uint64_t chunk = (n + 1023) / 1024;
    const uint64_t b = (uint64_t)threadIdx.x * chunk, e = b + chunk < n ? b + chunk : n;
    uint64_t last = 0;
    for (uint64_t i = b; i < e; ++i) if (a[i]) last = a[i];
    part[threadIdx.x] = last;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t run = 0;
        for (int t = 0; t < 1024; ++t) { const uint64_t v = part[t]; part[t] = run; if (v) run = v; }
    }
    __syncthreads();
    uint64_t run = part[threadIdx.x];
    for (uint64_t i = b; i < e; ++i) { const uint64_t v = a[i]; a[i] = run; if (v) run = v; }
}

// ---- passes 3 and 5 -------------------------------------------------------------------------------------------------
struct WalkArgs {
    const uint8_t* text; uint64_t n; bool aligned; int iupac;
    const uint64_t* entry;                 // per tile: FASTA kind of the last delimiter before it, FASTQ newlines before it
    uint64_t *rec, *seq, *qual;            // per tile: counts (pass 3 writes) / their exclusive prefix sums (pass 5 reads)
    uint8_t* out_seqs; uint64_t* out_limits; uint64_t limit_base;
    unsigned long long* g;
};

__device__ __forceinline__ void report(unsigned long long* g, uint64_t pos, uint32_t code)
{
    atomicMin(&g[G_ERR], (unsigned long long)((pos << 8) | code));
}

struct Run { uint64_t rec, seq, qual; };   // running totals (pass 3: this thread's; pass 5: global)

template <bool EMIT>
__device__ __forceinline__ void walk_fasta(const WalkArgs& a, const Chunk& c, uint64_t i0, uint64_t begin, bool in_id, Run& r)
{
    #pragma unroll
    for (int j = 0; j < IB; ++j) {
        if ((uint32_t)j >= c.m) break;
        const uint64_t i = i0 + j;
        const uint32_t ch = c.at(j);
        if (i < begin) continue;                           // skipUntil(iter, '>') in front of the first record
        if (in_id) { if (is_nl(ch)) in_id = false; }
        else if (ch == '>') { if (EMIT) a.out_limits[r.rec] = a.limit_base + r.seq; ++r.rec; in_id = true; }
        else if (!is_ws(ch)) {
            if (!EMIT && !in_alphabet(ch, a.iupac)) report(a.g, i, E_ALPHABET);
            if (EMIT) a.out_seqs[r.seq] = (uint8_t)ch;
            ++r.seq;
        }
    }
}

template <bool EMIT>
__device__ __forceinline__ void walk_fastq(const WalkArgs& a, const Chunk& c, uint64_t i0, uint64_t end, uint64_t line, uint32_t prev, Run& r)
{
    #pragma unroll
    for (int j = 0; j < IB; ++j) {
        if ((uint32_t)j >= c.m) break;
        const uint64_t i = i0 + j;
        const uint32_t ch = c.at(j);
        const uint32_t p = (uint32_t)line & 3u;
        const bool first = prev == '\n';
        if (p == 0) {
            if (first) {
                if (ch == '@') { if (EMIT) a.out_limits[r.rec] = a.limit_base + r.seq; ++r.rec; }
                else if (!EMIT) report(a.g, i, E_FQ_LAYOUT);
            }
        } else if (p == 1) {
            if (!is_ws(ch)) {
                if (!EMIT && !in_alphabet(ch, a.iupac)) report(a.g, i, E_ALPHABET);
                if (EMIT) a.out_seqs[r.seq] = (uint8_t)ch;
                ++r.seq;
            }
        } else if (p == 2) {
            if (first && ch != '+' && !EMIT) report(a.g, i, E_FQ_LAYOUT);
        } else if (!is_ws(ch)) ++r.qual;
        if (ch == '\n') {
            if (EMIT && p == 3 && r.seq != r.qual) report(a.g, i, E_FQ_QUAL);
            ++line;
        }
        if (EMIT && i + 1 == end && !((p == 3 || p == 2) && r.seq == r.qual)) report(a.g, i, p >= 2 ? E_FQ_QUAL : E_FQ_TRUNC);
        prev = ch;
    }
}

template <int FMT, bool EMIT>
__global__ void __launch_bounds__(IT) ingest_walk_kernel(const WalkArgs a)
{
    __shared__ uint64_t sm[IT / 32];
    const uint64_t begin = FMT == DYIBF_FMT_FASTA ? a.g[G_BEGIN] : 0;
    const uint64_t end = FMT == DYIBF_FMT_FASTA ? a.n : a.g[G_END];
    const uint64_t i0 = (uint64_t)blockIdx.x * TILE + (uint64_t)threadIdx.x * IB;
    const Chunk c = load_chunk(a.text, i0, end, a.aligned);
    uint64_t total;
    // state at this thread's first byte
    uint64_t mine = 0;
    #pragma unroll
    for (int j = 0; j < IB; ++j) {
        if ((uint32_t)j >= c.m) break;
        const uint32_t ch = c.at(j);
        if (FMT == DYIBF_FMT_FASTA) { if (ch == '>') mine = D_GT; else if (is_nl(ch)) mine = D_NL; }
        else mine += ch == '\n';
    }
    const uint64_t tile_entry = a.entry[blockIdx.x];
    uint64_t state;
    if (FMT == DYIBF_FMT_FASTA) { state = block_excl_scan(mine, OpLast(), sm, &total); if (!state) state = tile_entry; }
    else state = tile_entry + block_excl_scan(mine, OpSum(), sm, &total);
    const uint32_t prev = FMT == DYIBF_FMT_FASTQ ? (i0 == 0 || c.m == 0 ? (uint32_t)'\n' : (uint32_t)__ldg(a.text + i0 - 1)) : 0u;

    Run r; r.rec = r.seq = r.qual = 0;
    if (FMT == DYIBF_FMT_FASTA) walk_fasta<false>(a, c, i0, begin, state == D_GT, r);
    else                        walk_fastq<false>(a, c, i0, end, state, prev, r);
    const uint64_t packed = r.rec | (r.seq << 16) | (r.qual << 32);       // each at most 16 per thread, 4096 per tile
    const uint64_t excl = block_excl_scan(packed, OpSum(), sm, &total);
    if (!EMIT) {
        if (threadIdx.x == 0) {
            a.rec[blockIdx.x] = total & 0xffffu; a.seq[blockIdx.x] = (total >> 16) & 0xffffu; a.qual[blockIdx.x] = (total >> 32) & 0xffffu;
        }
        return;
    }
    r.rec  = a.rec[blockIdx.x]  + (excl & 0xffffu);
    r.seq  = a.seq[blockIdx.x]  + ((excl >> 16) & 0xffffu);
    r.qual = a.qual[blockIdx.x] + ((excl >> 32) & 0xffffu);
    if (FMT == DYIBF_FMT_FASTA) walk_fasta<true>(a, c, i0, begin, state == D_GT, r);
    else                        walk_fastq<true>(a, c, i0, end, state, prev, r);
}

__global__ void ingest_init_globals(unsigned long long* g)
{
    g[G_BEGIN] = ~0ull; g[G_END] = 0; g[G_ERR] = ~0ull; g[3] = 0;
}

__global__ void ingest_set_u64(uint64_t* p, uint64_t v) { *p = v; }

} // namespace

size_t ingest_workspace_bytes(uint64_t n_bytes)
{
    const uint64_t n_tiles = (n_bytes + TILE - 1) / TILE;
    return (size_t)((n_tiles + 1 + 3 * scan_workspace_elems(n_tiles) + G_N) * sizeof(uint64_t));
}

// Passes 1-4.  ws: ingest_workspace_bytes(n) bytes.  After the stream has drained, totals[0..2] = records, sequence
// characters, quality characters and totals[3..5] = first '>' offset (FASTA), effective end (FASTQ), first error word —
// all read back by the caller from `ws` through ingest_read_totals.
cudaError_t launch_ingest_count(const uint8_t* text, uint64_t n, int format, int iupac, uint64_t* ws, cudaStream_t st)
{
    const uint64_t n_tiles = (n + TILE - 1) / TILE;
    const size_t se = scan_workspace_elems(n_tiles);
    uint64_t* entry = ws; uint64_t* rec = entry + n_tiles + 1; uint64_t* seq = rec + se; uint64_t* qual = seq + se;
    unsigned long long* g = reinterpret_cast<unsigned long long*>(qual + se);
    const bool aligned = (reinterpret_cast<uintptr_t>(text) & 15u) == 0;
    ingest_init_globals<<<1, 1, 0, st>>>(g);
    WalkArgs a; a.text = text; a.n = n; a.aligned = aligned; a.iupac = iupac; a.entry = entry; a.rec = rec; a.seq = seq; a.qual = qual;
    a.out_seqs = nullptr; a.out_limits = nullptr; a.limit_base = 0; a.g = g;
    cudaError_t e;
    if (format == DYIBF_FMT_FASTA) {
        ingest_summary_kernel<DYIBF_FMT_FASTA><<<(unsigned)n_tiles, IT, 0, st>>>(text, n, aligned, entry, g);
        ingest_propagate_kernel<<<1, 1024, 0, st>>>(entry, n_tiles);
        ingest_walk_kernel<DYIBF_FMT_FASTA, false><<<(unsigned)n_tiles, IT, 0, st>>>(a);
    } else {
        ingest_summary_kernel<DYIBF_FMT_FASTQ><<<(unsigned)n_tiles, IT, 0, st>>>(text, n, aligned, entry, g);
        // entry has n_tiles + 1 elements; the scan's scratch is borrowed from `rec`, which pass 3 overwrites afterwards
        if ((e = cudaMemcpyAsync(rec, entry, n_tiles * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
        if ((e = launch_scan_u64(rec, n_tiles, st)) != cudaSuccess) return e;
        if ((e = cudaMemcpyAsync(entry, rec, (n_tiles + 1) * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
        ingest_walk_kernel<DYIBF_FMT_FASTQ, false><<<(unsigned)n_tiles, IT, 0, st>>>(a);
    }
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if ((e = launch_scan_u64(rec, n_tiles, st)) != cudaSuccess) return e;
    if ((e = launch_scan_u64(seq, n_tiles, st)) != cudaSuccess) return e;
    return launch_scan_u64(qual, n_tiles, st);
}

cudaError_t ingest_read_totals(const uint64_t* ws, uint64_t n, uint64_t totals[6], cudaStream_t st)
{
    const uint64_t n_tiles = (n + TILE - 1) / TILE;
    const size_t se = scan_workspace_elems(n_tiles);
    const uint64_t* rec = ws + n_tiles + 1; const uint64_t* seq = rec + se; const uint64_t* qual = seq + se; const uint64_t* g = qual + se;
    cudaError_t e;
    if ((e = cudaMemcpyAsync(&totals[0], rec + n_tiles, 8, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
    if ((e = cudaMemcpyAsync(&totals[1], seq + n_tiles, 8, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
    if ((e = cudaMemcpyAsync(&totals[2], qual + n_tiles, 8, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
    if ((e = cudaMemcpyAsync(&totals[3], g, 24, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
    return cudaStreamSynchronize(st);
}

// Pass 5: out_seqs[n_bases], out_limits[n_records + 1] (device), limits offset by limit_base.
cudaError_t launch_ingest_emit(const uint8_t* text, uint64_t n, int format, int iupac, uint64_t* ws, uint8_t* out_seqs,
                               uint64_t* out_limits, uint64_t limit_base, uint64_t n_records, uint64_t n_bases, cudaStream_t st)
{
    const uint64_t n_tiles = (n + TILE - 1) / TILE;
    const size_t se = scan_workspace_elems(n_tiles);
    WalkArgs a; a.text = text; a.n = n; a.aligned = (reinterpret_cast<uintptr_t>(text) & 15u) == 0; a.iupac = iupac;
    a.entry = ws; a.rec = ws + n_tiles + 1; a.seq = a.rec + se; a.qual = a.seq + se;
    a.out_seqs = out_seqs; a.out_limits = out_limits; a.limit_base = limit_base;
    a.g = reinterpret_cast<unsigned long long*>(a.qual + se);
    if (format == DYIBF_FMT_FASTA) ingest_walk_kernel<DYIBF_FMT_FASTA, true><<<(unsigned)n_tiles, IT, 0, st>>>(a);
    else                           ingest_walk_kernel<DYIBF_FMT_FASTQ, true><<<(unsigned)n_tiles, IT, 0, st>>>(a);
    ingest_set_u64<<<1, 1, 0, st>>>(out_limits + n_records, limit_base + n_bases);
    return cudaGetLastError();
}

} // namespace dyibf
