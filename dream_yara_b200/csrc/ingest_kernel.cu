// ingest_kernel.cu — FASTA / FASTQ text -> sequence characters + limits on the device (dyibf_parse_records).
//
// Replaces, for the step in front of the filter, SeqAn's sequential tokenizer as the reference drives it:
// readRecords(me.names, me.seqs, fileIn) in loadReads (src/mapper.h:451, src/store_seqs.h:175-188) and the readRecord
// loop of addFastaFile (src/d_bloom_filter.h:127-147); rules from seqan seq_io/fasta_fastq.h:268-282 (Fasta),
// :302-345 (Fastq), stream/tokenization.h:51-58 (IsInAlphabet), :145-149 (whitespace = ' ' '\t' '\n' '\r').
//
// Both formats are state machines whose state at a byte is a prefix function of the text, so the text is cut into tiles
// of 8 KiB (256 threads x 32 bytes) and parsed in streaming passes:
//   1 summary : per tile, what a later tile needs to know about it, and its counts under every state it can be entered in;
//               one read of the text
//   2 scan    : tile summaries -> state at the start of every tile                (n/8192 elements)
//   3 select  : per tile, records started / sequence characters / quality characters for the state it IS entered in
//   4 scan    : exclusive prefix sums of those counts over the tiles
//   5 emit    : validation, sequence characters to out_seqs, limits[r] at every record start;  second read
// (Round 1 counted in a pass of its own over the text — three reads.)
// Pass 1's counts per entry state — FASTQ: per class of LOCAL line number mod 4, the non-white-space characters and the lines opening with '@'; FASTA: the
// record / sequence counts for "entered outside an id", "entered inside an id" and "holds the first '>' of the text" — make
// pass 3 a selection over the tile table (ingest_select_fastq / _fasta); the validation sits in the emit pass (a sizing call
// therefore runs the emit pass into scratch).  FASTQ: 330 -> 405 GB/s of text.
// FASTA  — SeqAn: skip to the next '>', the rest of that line is the id, then every non-whitespace character up to the
//   next '>' (anywhere, not only at a line start) is sequence.  An id ends at a newline and a '>' inside an id is text, so
//   "inside an id" at byte i <=> the last of {'\n', '\r', '>'} before i is a '>': the tile summary is the kind of its last
//   such character, the scan propagates the last non-empty summary.  Characters before the first '>' are skipped.
// FASTQ  — the four-line layout ('@' id / sequence / '+' / qualities, one line each): the state is the number of '\n'
//   before the byte, mod 4; the tile summary is its newline count, the scan a prefix sum.  SeqAn also accepts sequences
//   and qualities wrapped over several lines and blank lines between records; those need the sequential reader (the
//   number of quality characters to consume is the length of the sequence read so far) and are handed to the general parser
//   (ingest_general.cu) when this tokeniser finds the layout violated: every '@' line must open a line whose number is 0 mod 4, every third
//   line must open with '+', and at every end of a quality line the running totals of sequence and quality characters
//   must agree — which is also SeqAn's own check (:341-344).  Trailing whitespace after the last record is ignored.
// Characters outside the target alphabet are SeqAn ParseErrors (reads: ACGTN, contigs: the 16 IUPAC letters; either
// case); here the smallest offending offset is reported through dyibf_last_error.
//
// Inside a tile nothing walks byte by byte: a thread turns its 32 bytes into 32-bit masks with byte-wise SIMD compares
// (newline, '>', '@', '+', whitespace, alphabet), derives "sequence character" / "quality character" / "record start" masks
// from them with one short loop over the delimiters it holds (usually none or one), counts with popc, and the emit pass
// stages the tile's sequence characters in shared memory — shifted so that 16-byte chunks line up with the output — and
// writes them as whole 16-byte stores.  (v1 walked the bytes with a four-way branch per byte: lanes of a warp sit in
// different lines of a record, so every branch diverged; 93 GB/s of text.  profiles/r01_ingest_*.)
#include "dyibf_internal.h"

#include <cstdio>

namespace dyibf {

namespace {

const uint32_t FULL = 0xffffffffu;
const int      IT = 256;                   // threads per tile
const int      IB = 32;                    // bytes per thread: one bit per byte in a 32-bit mask
const uint64_t TILE = (uint64_t)IT * IB;

enum { G_BEGIN = 0, G_END = 1, G_ERR = 2, G_N = 4 };     // device globals: first '>' / effective end / first error
enum { E_ALPHABET = 1, E_FQ_LAYOUT = 2, E_FQ_QUAL = 3, E_FQ_TRUNC = 4 };
enum { D_NONE = 0, D_NL = 1, D_GT = 2 };

// ---- 32 bytes of text in registers, and masks over them ----------------------------------------------------------
struct Chunk {
    uint32_t w[8];
    uint32_t valid;            // bit j: byte j exists (below the end of the text)
};

__device__ __forceinline__ Chunk load_chunk(const uint8_t* __restrict__ text, uint64_t i0, uint64_t end, bool aligned)
{
    Chunk c;
    #pragma unroll
    for (int q = 0; q < 8; ++q) c.w[q] = 0;
    const uint32_t m = i0 >= end ? 0u : (uint32_t)(end - i0 < (uint64_t)IB ? end - i0 : (uint64_t)IB);
    c.valid = m == 32 ? FULL : ((1u << m) - 1u);
    if (m == IB && aligned) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(text + i0));
        const uint4 b = __ldg(reinterpret_cast<const uint4*>(text + i0 + 16));
        c.w[0] = a.x; c.w[1] = a.y; c.w[2] = a.z; c.w[3] = a.w; c.w[4] = b.x; c.w[5] = b.y; c.w[6] = b.z; c.w[7] = b.w;
    } else {
        #pragma unroll
        for (int j = 0; j < IB; ++j)
            if ((uint32_t)j < m) c.w[j >> 2] |= (uint32_t)__ldg(text + i0 + j) << (8 * (j & 3));
    }
    return c;
}

// 0x80 in every byte of x that equals ch, 0 elsewhere (exact: no borrow between bytes; __vcmpeq4 is emulated with more)
__device__ __forceinline__ uint32_t eq4(uint32_t x, uint32_t ch)
{
    const uint32_t d = x ^ (ch * 0x01010101u);                        // zero byte <=> equal
    return ~(((d & 0x7f7f7f7fu) + 0x7f7f7f7fu) | d) & 0x80808080u;
}
// the top bit of each of the four bytes (nothing else set) -> a 4-bit mask (byte 0 -> bit 0)
__device__ __forceinline__ uint32_t movemask4(uint32_t x) { return ((x >> 7) * 0x01020408u) >> 24; }

struct ByteTest {   // one functor per byte class: word -> 0x80 / 0x00 per byte
    struct Char  { uint32_t ch; __device__ __forceinline__ uint32_t operator()(uint32_t x) const { return eq4(x, ch); } };
    struct NlCr  { __device__ __forceinline__ uint32_t operator()(uint32_t x) const { return eq4(x, '\n') | eq4(x, '\r'); } };
    struct White { __device__ __forceinline__ uint32_t operator()(uint32_t x) const { return eq4(x, ' ') | eq4(x, '\t') | eq4(x, '\n') | eq4(x, '\r'); } };
    // IsInAlphabet<Dna5Q> / <Iupac>: clearing bit 5 maps a lower-case letter onto its upper-case one and nothing else onto a letter
    struct Alpha {
        int iupac;
        __device__ __forceinline__ uint32_t operator()(uint32_t x) const {
            const uint32_t u = x & 0xdfdfdfdfu;
            uint32_t ok = eq4(u, 'A') | eq4(u, 'C') | eq4(u, 'G') | eq4(u, 'T') | eq4(u, 'N');
            if (iupac) ok |= eq4(u, 'U') | eq4(u, 'W') | eq4(u, 'Y') | eq4(u, 'M') | eq4(u, 'H') | eq4(u, 'K') | eq4(u, 'R') |
                             eq4(u, 'D') | eq4(u, 'S') | eq4(u, 'B') | eq4(u, 'V');
            return ok;
        }
    };
};

template <typename Test>
__device__ __forceinline__ uint32_t mask_of(const Chunk& c, Test t)
{
    uint32_t m = 0;
    #pragma unroll
    for (int q = 0; q < 8; ++q) m |= movemask4(t(c.w[q])) << (4 * q);
    return m;
}

// bits [lo, hi) of a 32-bit mask, 0 <= lo <= hi <= 32
__device__ __forceinline__ uint32_t bit_range(uint32_t lo, uint32_t hi)
{
    const uint32_t upto_hi = hi >= 32 ? FULL : ((1u << hi) - 1u);
    const uint32_t upto_lo = lo >= 32 ? FULL : ((1u << lo) - 1u);
    return upto_hi & ~upto_lo;
}

struct OpSum  { template <typename T> __device__ __forceinline__ T operator()(T a, T b) const { return a + b; } };
struct OpLast { template <typename T> __device__ __forceinline__ T operator()(T a, T b) const { return b ? b : a; } };   // a earlier

// exclusive scan over the 256 threads of a block (identity 0); *total = combination of all
template <typename T, typename Op>
__device__ __forceinline__ T block_excl_scan(T v, Op op, T* sm /* [IT/32] */, T* total)
{
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T x = v;
    #pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const T y = __shfl_up_sync(FULL, x, d);
        if (lane >= (uint32_t)d) x = op(y, x);
    }
    __syncthreads();                                   // sm may still be read by a previous scan
    if (lane == 31) sm[warp] = x;
    __syncthreads();
    if (warp == 0) {
        T w = lane < IT / 32 ? sm[lane] : (T)0;
        #pragma unroll
        for (int d = 1; d < IT / 32; d <<= 1) {
            const T y = __shfl_up_sync(FULL, w, d);
            if (lane >= (uint32_t)d) w = op(y, w);
        }
        if (lane < IT / 32) sm[lane] = w;
    }
    __syncthreads();
    T excl = __shfl_up_sync(FULL, x, 1);
    if (lane == 0) excl = 0;
    const T prefix = warp ? sm[warp - 1] : (T)0;
    *total = sm[IT / 32 - 1];
    return op(prefix, excl);
}

// ---- what a thread's bytes are, once the state at its first byte is known ------------------------------------------
struct Marks {
    uint32_t seq;        // sequence characters
    uint32_t qual;       // FASTQ: quality characters
    uint32_t rec;        // record starts ('>' outside an id; '@' opening a line 0 mod 4)
    uint32_t bad;        // FASTQ: a line 0 mod 4 that does not open with '@', a line 2 mod 4 that does not open with '+'
    uint32_t qual_eol;   // FASTQ: the '\n' that ends a quality line
    uint32_t late;       // FASTQ: bytes of lines 2 and 3 mod 4 (where a text may end)
};

// FASTA.  gt / nl: '>' and newline masks; in_id: state at byte 0; bytes before `begin` (the first '>') are skipped.
__device__ __forceinline__ Marks mark_fasta(const Chunk& c, uint32_t gt, uint32_t nl, uint32_t ws, bool in_id, uint64_t i0, uint64_t begin)
{
    uint32_t id = 0, rec = 0, cur = 0, rem = (gt | nl) & c.valid;
    bool st = in_id;
    for (;;) {                                               // one round per delimiter held: usually none or one
        const uint32_t nxt = rem ? (uint32_t)__ffs(rem) - 1u : 32u;
        if (st) id |= bit_range(cur, nxt);
        if (!rem) break;
        const uint32_t bit = 1u << nxt;
        if (gt & bit) { if (st) id |= bit; else rec |= bit; st = true; }   // a '>' inside an id is text
        else          { if (st) id |= bit; st = false; }                   // the newline that ends the id
        rem &= rem - 1; cur = nxt + 1;
    }
    const uint32_t after = i0 >= begin ? FULL : (begin - i0 >= 32 ? 0u : ~((1u << (uint32_t)(begin - i0)) - 1u));
    Marks m; m.qual = m.bad = m.qual_eol = m.late = 0;
    m.rec = rec;
    m.seq = c.valid & after & ~id & ~gt & ~ws;
    return m;
}

// FASTQ, four-line layout.  nl: '\n' mask; line0: newlines before byte 0; prev_nl: byte -1 is a '\n' (or the text starts here)
__device__ __forceinline__ Marks mark_fastq(const Chunk& c, uint32_t nl, uint32_t ws, uint32_t at, uint32_t plus, uint64_t line0, bool prev_nl)
{
    uint32_t p0 = 0, p1 = 0, p2 = 0, p3 = 0, cur = 0, rem = nl & c.valid;
    uint32_t p = (uint32_t)line0 & 3u;
    for (;;) {                                               // one round per line end held
        const uint32_t nxt = rem ? (uint32_t)__ffs(rem) : 32u;       // one past the newline: it belongs to the line it ends
        const uint32_t seg = bit_range(cur, nxt);
        p0 |= p == 0 ? seg : 0u; p1 |= p == 1 ? seg : 0u; p2 |= p == 2 ? seg : 0u; p3 |= p == 3 ? seg : 0u;
        if (!rem) break;
        rem &= rem - 1; cur = nxt; p = (p + 1) & 3u;
    }
    const uint32_t first = (nl << 1) | (prev_nl ? 1u : 0u);
    Marks m;
    m.seq = p1 & ~ws & c.valid;
    m.qual = p3 & ~ws & c.valid;
    m.rec = p0 & first & at;
    m.bad = ((p0 & first & ~at) | (p2 & first & ~plus)) & c.valid;
    m.qual_eol = p3 & nl & c.valid;
    m.late = (p2 | p3) & c.valid;
    return m;
}

__device__ __forceinline__ void report(unsigned long long* g, uint64_t pos, uint32_t code)
{
    atomicMin(&g[G_ERR], (unsigned long long)((pos << 8) | code));
}

// ---- pass 1 ----------------------------------------------------------------------------------------------------------
// FASTQ: besides its newline count a tile reports, per class j of LOCAL line number mod 4 (lines counted from the tile's first
// byte), the non-white-space characters (cls[2 t], 16 bits per class) and the lines that open with '@' (cls[2 t + 1]).  Once
// the scan has told how many lines precede the tile, the class that holds sequences / qualities / record starts is known and
// the tile's counts are picked from these (ingest_select_fastq) — the text is not read a second time for counting.
template <int FMT>
__global__ void __launch_bounds__(IT) ingest_summary_kernel(const uint8_t* __restrict__ text, uint64_t n, bool aligned,
                                                            uint64_t* __restrict__ summary, unsigned long long* __restrict__ g,
                                                            uint64_t* __restrict__ cls)
{
    __shared__ uint32_t sm[IT / 32];
    __shared__ uint64_t sm64[IT / 32];
    const uint64_t i0 = (uint64_t)blockIdx.x * TILE + (uint64_t)threadIdx.x * IB;
    const Chunk c = load_chunk(text, i0, n, aligned);
    uint32_t total;
    if (FMT == DYIBF_FMT_FASTA) {
        const uint32_t gt = mask_of(c, ByteTest::Char{'>'}), nl = mask_of(c, ByteTest::NlCr());
        const uint32_t mine = (gt | nl) ? (gt > nl ? D_GT : D_NL) : D_NONE;      // the higher mask holds the later byte
        const uint32_t any = __ballot_sync(FULL, gt != 0);                         // the warp's first '>' is its lowest lane's
        if (any && (threadIdx.x & 31) == (uint32_t)(__ffs(any) - 1)) atomicMin(&g[G_BEGIN], (unsigned long long)(i0 + __ffs(gt) - 1));
        const uint32_t state = block_excl_scan<uint32_t>(mine, OpLast(), sm, &total);
        // Counts under the three views a tile can be in: A — it is entered outside an id; B — it is entered inside an id;
        // C — it holds the first '>' of the text, everything before that '>' is skipped.  Threads behind a delimiter of the tile
        // know their state; only the ones before the tile's first delimiter differ between A and B.
        __shared__ unsigned long long tile_first_gt;
        if (threadIdx.x == 0) tile_first_gt = ~0ull;
        __syncthreads();
        if (any && (threadIdx.x & 31) == (uint32_t)(__ffs(any) - 1)) atomicMin(&tile_first_gt, (unsigned long long)(i0 + __ffs(gt) - 1));
        __syncthreads();
        const uint64_t fg = tile_first_gt;
        const uint32_t ws = mask_of(c, ByteTest::White());
        const Marks ma = mark_fasta(c, gt, nl, ws, state == D_NONE ? false : state == D_GT, i0, 0);
        const Marks mb = state == D_NONE ? mark_fasta(c, gt, nl, ws, true, i0, 0) : ma;
        const uint32_t after = fg == ~0ull || fg >= i0 + 32 ? 0u : (fg <= i0 ? FULL : ~((1u << (uint32_t)(fg - i0)) - 1u));
        const uint64_t seqs = (uint64_t)__popc(ma.seq) | ((uint64_t)__popc(mb.seq) << 16) | ((uint64_t)__popc(ma.seq & after) << 32);
        const uint64_t recs = (uint64_t)__popc(ma.rec) | ((uint64_t)__popc(mb.rec) << 16) | ((uint64_t)__popc(ma.rec & after) << 32);
        uint64_t t_seqs, t_recs;
        block_excl_scan<uint64_t>(seqs, OpSum(), sm64, &t_seqs);         // each field at most 8192 per tile
        block_excl_scan<uint64_t>(recs, OpSum(), sm64, &t_recs);
        if (threadIdx.x == 0) { cls[2 * (uint64_t)blockIdx.x] = t_seqs; cls[2 * (uint64_t)blockIdx.x + 1] = t_recs; }
    } else {
        const uint32_t nl = mask_of(c, ByteTest::Char{'\n'}), graph = c.valid & ~mask_of(c, ByteTest::White());
        const uint64_t e = graph ? i0 + 32 - __clz(graph) : 0;                     // one past this thread's last non-whitespace byte
        uint64_t emax = e;
        #pragma unroll
        for (int d = 16; d; d >>= 1) { const uint64_t o = __shfl_xor_sync(FULL, emax, d); emax = o > emax ? o : emax; }
        if ((threadIdx.x & 31) == 0 && emax) atomicMax(&g[G_END], (unsigned long long)emax);
        const uint32_t local_line = block_excl_scan<uint32_t>(__popc(nl), OpSum(), sm, &total);
        // the thread's bytes by local line class: mark_fastq with the local line number as if the tile opened a record
        const uint32_t ws = mask_of(c, ByteTest::White());
        const bool prev_nl = i0 == 0 || (c.valid && __ldg(text + i0 - 1) == '\n');
        uint32_t pc[4] = {0, 0, 0, 0}, cur = 0, rem = nl & c.valid, p = local_line & 3u;
        for (;;) {
            const uint32_t nxt = rem ? (uint32_t)__ffs(rem) : 32u;
            const uint32_t seg = bit_range(cur, nxt);
            pc[0] |= p == 0 ? seg : 0u; pc[1] |= p == 1 ? seg : 0u; pc[2] |= p == 2 ? seg : 0u; pc[3] |= p == 3 ? seg : 0u;
            if (!rem) break;
            rem &= rem - 1; cur = nxt; p = (p + 1) & 3u;
        }
        const uint32_t first = (nl << 1) | (prev_nl ? 1u : 0u), at = mask_of(c, ByteTest::Char{'@'});
        uint64_t nonws = 0, ats = 0;
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
            nonws |= (uint64_t)__popc(pc[j] & ~ws & c.valid) << (16 * j);
            ats   |= (uint64_t)__popc(pc[j] & first & at & c.valid) << (16 * j);
        }
        uint64_t t_nonws, t_ats;
        block_excl_scan<uint64_t>(nonws, OpSum(), sm64, &t_nonws);       // each field at most 8192 per tile
        block_excl_scan<uint64_t>(ats, OpSum(), sm64, &t_ats);
        if (threadIdx.x == 0) { cls[2 * (uint64_t)blockIdx.x] = t_nonws; cls[2 * (uint64_t)blockIdx.x + 1] = t_ats; }
    }
    if (threadIdx.x == 0) summary[blockIdx.x] = total;
}

// FASTA: the tile's record / sequence counts from its three views, the state it is entered in and the first '>' of the text
__global__ void __launch_bounds__(256) ingest_select_fasta(const uint64_t* __restrict__ entry, const uint64_t* __restrict__ cls, uint64_t n_tiles,
                                                           const unsigned long long* __restrict__ g, uint64_t* __restrict__ rec,
                                                           uint64_t* __restrict__ seq, uint64_t* __restrict__ qual)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tiles) return;
    const uint64_t begin = g[G_BEGIN], tb = begin == ~0ull ? ~0ull : begin / TILE;
    const int view = t < tb ? -1 : (t == tb ? 2 : (entry[t] == D_GT ? 1 : 0));
    rec[t] = view < 0 ? 0 : (cls[2 * t + 1] >> (16 * view)) & 0xffffu;
    seq[t] = view < 0 ? 0 : (cls[2 * t] >> (16 * view)) & 0xffffu;
    qual[t] = 0;
}

// FASTQ: the tile's record / sequence / quality counts from its class counts and the lines before it
__global__ void __launch_bounds__(256) ingest_select_fastq(const uint64_t* __restrict__ entry, const uint64_t* __restrict__ cls, uint64_t n_tiles,
                                                           uint64_t* __restrict__ rec, uint64_t* __restrict__ seq, uint64_t* __restrict__ qual)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tiles) return;
    const uint32_t d = (uint32_t)entry[t] & 3u;                          // global line number of the tile's local line 0, mod 4
    const uint64_t nonws = cls[2 * t], ats = cls[2 * t + 1];
    // local class j holds global lines (j + d) mod 4: records open global lines 0, sequences are lines 1, qualities lines 3
    rec[t]  = (ats   >> (16 * ((0u - d) & 3u))) & 0xffffu;
    seq[t]  = (nonws >> (16 * ((1u - d) & 3u))) & 0xffffu;
    qual[t] = (nonws >> (16 * ((3u - d) & 3u))) & 0xffffu;
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
    uint64_t* out_rec_pos;                 // optional: text offset of every record's start character
    unsigned long long* g;
};

template <int FMT, bool EMIT>
__global__ void __launch_bounds__(IT) ingest_walk_kernel(const WalkArgs a)
{
    __shared__ uint64_t sm[IT / 32];
    __shared__ __align__(16) uint8_t stage[EMIT ? TILE + 16 : 16];
    const uint64_t begin = FMT == DYIBF_FMT_FASTA ? a.g[G_BEGIN] : 0;
    const uint64_t end = FMT == DYIBF_FMT_FASTA ? a.n : a.g[G_END];
    const uint64_t i0 = (uint64_t)blockIdx.x * TILE + (uint64_t)threadIdx.x * IB;
    const Chunk c = load_chunk(a.text, i0, end, a.aligned);
    const uint32_t ws = mask_of(c, ByteTest::White());

    // state at this thread's first byte, then what its bytes are
    Marks m;
    uint64_t total;
    if (FMT == DYIBF_FMT_FASTA) {
        const uint32_t gt = mask_of(c, ByteTest::Char{'>'}), nl = mask_of(c, ByteTest::NlCr());
        const uint64_t mine = (gt | nl) ? (gt > nl ? D_GT : D_NL) : D_NONE;
        uint64_t state = block_excl_scan<uint64_t>(mine, OpLast(), sm, &total);
        if (!state) state = a.entry[blockIdx.x];
        m = mark_fasta(c, gt, nl, ws, state == D_GT, i0, begin);
    } else {
        const uint32_t nl = mask_of(c, ByteTest::Char{'\n'});
        const uint64_t line0 = a.entry[blockIdx.x] + block_excl_scan<uint64_t>(__popc(nl), OpSum(), sm, &total);
        const bool prev_nl = i0 == 0 || (c.valid && __ldg(a.text + i0 - 1) == '\n');
        m = mark_fastq(c, nl, ws, mask_of(c, ByteTest::Char{'@'}), mask_of(c, ByteTest::Char{'+'}), line0, prev_nl);
    }
    if (EMIT) {                                             // SeqAn's ParseErrors, reported once: the emit pass is the one that looks at every byte's role
        const uint32_t foreign = m.seq & ~mask_of(c, ByteTest::Alpha{a.iupac});
        if (foreign) report(a.g, i0 + __ffs(foreign) - 1, E_ALPHABET);
        if (m.bad)   report(a.g, i0 + __ffs(m.bad) - 1, E_FQ_LAYOUT);
    }
    const uint64_t packed = (uint64_t)__popc(m.rec) | ((uint64_t)__popc(m.seq) << 16) | ((uint64_t)__popc(m.qual) << 32);
    const uint64_t excl = block_excl_scan<uint64_t>(packed, OpSum(), sm, &total);   // each field at most 8192 per tile
    if (!EMIT) {
        if (threadIdx.x == 0) {
            a.rec[blockIdx.x] = total & 0xffffu; a.seq[blockIdx.x] = (total >> 16) & 0xffffu; a.qual[blockIdx.x] = (total >> 32) & 0xffffu;
        }
        return;
    }
    const uint64_t tile_seq = a.seq[blockIdx.x];
    const uint32_t my_seq = (uint32_t)(excl >> 16) & 0xffffu, tile_cnt = (uint32_t)(total >> 16) & 0xffffu;
    uint64_t rec_at = a.rec[blockIdx.x] + (excl & 0xffffu);
    const uint64_t seq_at = tile_seq + my_seq, qual_at = a.qual[blockIdx.x] + ((excl >> 32) & 0xffffu);

    // limits[r] at every record start
    for (uint32_t r = m.rec; r; r &= r - 1) {
        const uint32_t below = (1u << (__ffs(r) - 1)) - 1u;
        if (a.out_rec_pos) a.out_rec_pos[rec_at] = i0 + (uint32_t)(__ffs(r) - 1);
        a.out_limits[rec_at++] = a.limit_base + seq_at + __popc(m.seq & below);
    }
    if (FMT == DYIBF_FMT_FASTQ) {
        // SeqAn consumes exactly length(seq) quality characters: at the end of every quality line the totals must agree
        for (uint32_t q = m.qual_eol; q; q &= q - 1) {
            const uint32_t j = __ffs(q) - 1, below = (1u << j) - 1u;
            if (seq_at + __popc(m.seq & below) != qual_at + __popc(m.qual & below)) report(a.g, i0 + j, E_FQ_QUAL);
        }
        if (c.valid && i0 + 32 - __clz(c.valid) == end) {   // this thread holds the last byte of the text
            const uint32_t last = 1u << (31 - __clz(c.valid));
            if (!(m.late & last) || seq_at + __popc(m.seq) != qual_at + __popc(m.qual))
                report(a.g, end - 1, (m.late & last) ? E_FQ_QUAL : E_FQ_TRUNC);
        }
    }
    // the tile's sequence characters: staged so that 16-byte chunks of shared memory are 16-byte chunks of the output
    uint8_t* dst = a.out_seqs + tile_seq;
    const uint32_t shift = (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u);
    uint32_t k = shift + my_seq;
    #pragma unroll
    for (int j = 0; j < IB; ++j) {
        const bool on = (m.seq >> j) & 1u;
        if (on) stage[k] = (uint8_t)(c.w[j >> 2] >> (8 * (j & 3)));
        k += on;
    }
    __syncthreads();
    const uint32_t lo = shift, hi = shift + tile_cnt;
    uint8_t* dst_al = dst - shift;
    for (uint32_t b = threadIdx.x * 16u; b < hi; b += IT * 16u) {
        if (b >= lo && b + 16 <= hi) *reinterpret_cast<uint4*>(dst_al + b) = *reinterpret_cast<const uint4*>(stage + b);
        else
            for (uint32_t x = b < lo ? lo : b; x < b + 16 && x < hi; ++x) dst_al[x] = stage[x];
    }
}

__global__ void ingest_init_globals(unsigned long long* g)
{
    g[G_BEGIN] = ~0ull; g[G_END] = 0; g[G_ERR] = ~0ull; g[3] = 0;
}

__global__ void ingest_set_u64(uint64_t* p, uint64_t v) { *p = v; }

// longest sequence of a limits array (n + 1 entries)
__global__ void __launch_bounds__(256) ingest_max_len_kernel(const uint64_t* __restrict__ limits, uint64_t n, unsigned long long* __restrict__ out)
{
    uint64_t m = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t l = limits[i + 1] - limits[i];
        m = l > m ? l : m;
    }
    #pragma unroll
    for (int d = 16; d; d >>= 1) { const uint64_t o = __shfl_xor_sync(FULL, m, d); m = o > m ? o : m; }
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, (unsigned long long)m);
}

} // namespace

size_t ingest_workspace_bytes(uint64_t n_bytes)
{
    const uint64_t n_tiles = (n_bytes + TILE - 1) / TILE;
    return (size_t)((n_tiles + 1 + 3 * scan_workspace_elems(n_tiles) + G_N + 2 * n_tiles) * sizeof(uint64_t));
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
    a.out_seqs = nullptr; a.out_limits = nullptr; a.limit_base = 0; a.out_rec_pos = nullptr; a.g = g;
    cudaError_t e;
    if (format == DYIBF_FMT_FASTA) {
        uint64_t* cls = reinterpret_cast<uint64_t*>(g + G_N);
        ingest_summary_kernel<DYIBF_FMT_FASTA><<<(unsigned)n_tiles, IT, 0, st>>>(text, n, aligned, entry, g, cls);
        ingest_propagate_kernel<<<1, 1024, 0, st>>>(entry, n_tiles);
        // no second read of the text for counting: the tile's counts are picked from its three views
        ingest_select_fasta<<<(unsigned)((n_tiles + 255) / 256), 256, 0, st>>>(entry, cls, n_tiles, g, rec, seq, qual);
    } else {
        uint64_t* cls = reinterpret_cast<uint64_t*>(g + G_N);
        ingest_summary_kernel<DYIBF_FMT_FASTQ><<<(unsigned)n_tiles, IT, 0, st>>>(text, n, aligned, entry, g, cls);
        // entry has n_tiles + 1 elements; the scan's scratch is borrowed from `rec`, which the selection overwrites afterwards
        if ((e = cudaMemcpyAsync(rec, entry, n_tiles * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
        if ((e = launch_scan_u64(rec, n_tiles, st)) != cudaSuccess) return e;
        if ((e = cudaMemcpyAsync(entry, rec, (n_tiles + 1) * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
        // no second read of the text for counting: the tile's counts are picked from its class counts
        ingest_select_fastq<<<(unsigned)((n_tiles + 255) / 256), 256, 0, st>>>(entry, cls, n_tiles, rec, seq, qual);
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
                               uint64_t* out_limits, uint64_t limit_base, uint64_t n_records, uint64_t n_bases, cudaStream_t st,
                               uint64_t* out_rec_pos)
{
    const uint64_t n_tiles = (n + TILE - 1) / TILE;
    const size_t se = scan_workspace_elems(n_tiles);
    WalkArgs a; a.text = text; a.n = n; a.aligned = (reinterpret_cast<uintptr_t>(text) & 15u) == 0; a.iupac = iupac;
    a.entry = ws; a.rec = ws + n_tiles + 1; a.seq = a.rec + se; a.qual = a.seq + se;
    a.out_seqs = out_seqs; a.out_limits = out_limits; a.limit_base = limit_base; a.out_rec_pos = out_rec_pos;
    a.g = reinterpret_cast<unsigned long long*>(a.qual + se);
    if (format == DYIBF_FMT_FASTA) ingest_walk_kernel<DYIBF_FMT_FASTA, true><<<(unsigned)n_tiles, IT, 0, st>>>(a);
    else                           ingest_walk_kernel<DYIBF_FMT_FASTQ, true><<<(unsigned)n_tiles, IT, 0, st>>>(a);
    ingest_set_u64<<<1, 1, 0, st>>>(out_limits + n_records, limit_base + n_bases);
    return cudaGetLastError();
}

// *out (device, zeroed by the caller) = max over i of limits[i+1] - limits[i]
cudaError_t launch_max_len(const uint64_t* limits, uint64_t n, unsigned long long* out, cudaStream_t st)
{
    if (n == 0) return cudaSuccess;
    uint64_t blocks = (n + 255) / 256;
    if (blocks > sm_count_current() * 8) blocks = sm_count_current() * 8;
    ingest_max_len_kernel<<<(unsigned)blocks, 256, 0, st>>>(limits, n, out);
    return cudaGetLastError();
}

} // namespace dyibf
