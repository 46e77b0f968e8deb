// K9 -- on-device PNG decode: the DAQ's frames are 8-bit greyscale PNG files
// (AutoBubStart3.cpp:227-228 "cam%d_image%u.png"; cv::imread/imdecode in RawParser.cpp:39, ZipParser.cpp:222), and
// inflating them on the host cores is what bounds the reference-style ingest (~15 ms per frame per core).
//
//   k9_inflate_kernel   one warp per frame: RFC 1951 inflate of the frame's zlib stream (the concatenated IDAT
//                       payloads) into a scratch of rows x (cols+1) filtered scanlines.  All 32 lanes run the same
//                       control flow on warp-uniform values (bit buffer, Huffman tables in shared memory); lane 0
//                       stores literals, the whole warp copies LZ77 matches and builds the decoding tables.
//   k9_unfilter_kernel  one warp per frame: PNG filter reconstruction (None/Sub/Up/Average/Paeth, 1 byte per pixel)
//                       row by row through shared memory, written straight into the frame of the event slot.
// A batch of events is ~1300 frames, i.e. ~1300 warps in flight: the decode is latency-bound per stream and parallel
// across streams.  The compressed bytes travel H2D instead of the pixels (about half the volume).
// Lossless: the result is compared byte for byte with cv2.imdecode (tests/test_gpu_png.py).
#include "abub_internal.cuh"

namespace {

constexpr int TB_LIT = 10;                 // fast-table bits, literal/length codes
constexpr int TB_DIST = 8;                 // fast-table bits, distance codes
constexpr int K9_WARPS = 2;                // warps (frames) per CTA

struct PngDev {                            // per frame
    const uint8_t* z;                      // zlib stream (device)
    uint32_t zlen;
    uint8_t* filt;                         // rows * (cols + 1) filtered scanlines
    uint8_t* dst;                          // rows * cols output frame
};

struct WarpTables {
    uint16_t lit[1 << TB_LIT];             // (symbol << 4) | code length, 0 = not in the fast table
    uint16_t dist[1 << TB_DIST];
    uint16_t lit_sym[288], dist_sym[32];   // symbols in canonical order (slow path, codes longer than the table)
    uint16_t lit_cnt[16], dist_cnt[16];
    uint8_t len[320];                      // code lengths being read
};

// Bit reader shared by the warp.  The compressed stream is read one 128-byte line at a time: lane i holds word i of
// the current line (one coalesced load per 1024 bits) and of the next line (issued a whole line ahead, so its latency
// is never waited for); a refill broadcasts the next word with one SHFL.  All fields are warp-uniform except the two
// line registers.
struct Bits {
    const uint32_t* w;      // stream as 32-bit words (the frame region is 16-byte aligned)
    uint32_t nwords;        // words that may be read (stream length rounded up)
    uint32_t line;          // index of the current 32-word line
    int widx;               // next word of the line to take
    uint32_t cur, nxt;      // this lane's word of the current / next line
    uint32_t pre;           // word widx of the current line, already broadcast (keeps the SHFL off the refill's critical path)
    uint64_t buf;
    int cnt;
    bool overrun;
};

__device__ __forceinline__ uint32_t bits_load(const Bits& b, uint32_t line, int lane) {
    const uint32_t i = line * 32u + (uint32_t)lane;
    return i < b.nwords ? __ldg(b.w + i) : 0u;
}
__device__ __forceinline__ void bits_init(Bits& b, const uint8_t* p, uint32_t nbytes, uint32_t bytepos, int lane) {
    b.w = reinterpret_cast<const uint32_t*>(p);
    b.nwords = (nbytes + 3u) >> 2;
    const uint32_t wpos = bytepos >> 2;
    b.line = wpos >> 5;
    b.widx = (int)(wpos & 31u);
    b.cur = bits_load(b, b.line, lane);
    b.nxt = bits_load(b, b.line + 1, lane);
    b.pre = __shfl_sync(0xffffffffu, b.cur, b.widx);
    b.buf = 0;
    b.cnt = 0;
    b.overrun = false;
}
__device__ __forceinline__ void refill(Bits& b, int lane) {     // afterwards at least 33 bits are present
    if (b.cnt <= 32) {
        b.buf |= (uint64_t)b.pre << b.cnt;
        b.cnt += 32;
        if (++b.widx == 32) {
            b.widx = 0;
            b.line++;
            b.cur = b.nxt;
            b.nxt = bits_load(b, b.line + 1, lane);
            if (b.line * 32u > b.nwords + 32u) b.overrun = true;
        }
        b.pre = __shfl_sync(0xffffffffu, b.cur, b.widx);
    }
}
// byte offset in the stream of the first bit not yet consumed (call only when cnt is a multiple of 8)
__device__ __forceinline__ uint32_t bits_bytepos(const Bits& b) { return (b.line * 32u + (uint32_t)b.widx) * 4u - (uint32_t)(b.cnt >> 3); }
__device__ __forceinline__ uint32_t take(Bits& b, int n) {  // n <= 32, bits must be present
    const uint32_t v = (uint32_t)(b.buf & ((1ull << n) - 1ull));
    b.buf >>= n;
    b.cnt -= n;
    return v;
}

__device__ __forceinline__ uint32_t rev_bits(uint32_t c, int n) { return __brev(c) >> (32 - n); }

// Canonical Huffman tables from code lengths len[0..n): fast table of `tb` bits + count/symbol arrays.
// Returns false on an over-subscribed set.  Cooperative: every lane calls it.
__device__ bool build_tables(const uint8_t* len, int n, uint16_t* fast, int tb, uint16_t* cnt, uint16_t* sym, int lane) {
    for (int i = lane; i < (1 << tb); i += 32) fast[i] = 0;
    __syncwarp();
    bool ok = true;
    if (lane == 0) {
        uint16_t offs[16], next[16];
        for (int l = 0; l < 16; l++) cnt[l] = 0;
        for (int s = 0; s < n; s++) cnt[len[s]]++;
        int left = 1;
        for (int l = 1; l < 16; l++) {
            left <<= 1;
            left -= cnt[l];
            if (left < 0) ok = false;
        }
        offs[1] = 0;
        for (int l = 1; l < 15; l++) offs[l + 1] = offs[l] + cnt[l];
        uint32_t code = 0;
        next[0] = 0;
        for (int l = 1; l < 16; l++) {
            code = (code + cnt[l - 1] * (l > 1)) << 1;
            next[l] = (uint16_t)code;
        }
        // canonical first codes: code(l) = (code(l-1) + count(l-1)) << 1 with count(0) = 0
        for (int s = 0; s < n && ok; s++) {
            const int l = len[s];
            if (!l) continue;
            sym[offs[l]++] = (uint16_t)s;
            const uint32_t c = next[l]++;
            if (l <= tb) {
                const uint32_t r = rev_bits(c, l);
                for (uint32_t k = r; k < (1u << tb); k += (1u << l)) fast[k] = (uint16_t)((s << 4) | l);
            }
        }
    }
    ok = __shfl_sync(0xffffffffu, ok, 0);
    __syncwarp();
    return ok;
}

// Decodes one symbol (bits already refilled: at least 15 present or the input is at its end).
__device__ __forceinline__ int decode_sym(Bits& b, const uint16_t* fast, int tb, const uint16_t* cnt, const uint16_t* sym) {
    const uint32_t e = fast[(uint32_t)b.buf & ((1u << tb) - 1u)];
    if (e & 15u) {
        const int l = e & 15u;
        b.buf >>= l;
        b.cnt -= l;
        return (int)(e >> 4);
    }
    // slow path: canonical decode bit by bit (Mark Adler's puff scheme)
    int code = 0, first = 0, index = 0;
    uint64_t bits = b.buf;
    for (int l = 1; l <= 15; l++) {
        code |= (int)(bits & 1);
        bits >>= 1;
        const int c = cnt[l];
        if (code - c < first) {
            b.buf >>= l;
            b.cnt -= l;
            return sym[index + (code - first)];
        }
        index += c;
        first += c;
        first <<= 1;
        code <<= 1;
    }
    return -1;
}

__constant__ uint16_t c_len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
__constant__ uint8_t c_len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
__constant__ uint16_t c_dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
__constant__ uint8_t c_dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
__constant__ uint8_t c_clen_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

// status: 0 ok, 1 bad header, 2 bad block, 3 bad code, 4 output size mismatch, 5 input overrun
__global__ void __launch_bounds__(32 * K9_WARPS) k9_inflate_kernel(const PngDev* __restrict__ jobs, int njobs, uint32_t out_len,
                                                                   int* __restrict__ status) {
    __shared__ WarpTables tabs[K9_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int job = blockIdx.x * K9_WARPS + warp;
    if (job >= njobs) return;
    WarpTables& T = tabs[warp];
    const PngDev J = jobs[job];
    uint8_t* const out = J.filt;
    Bits b;
    bits_init(b, J.z, J.zlen, 0, lane);
    int st = 0;
    uint32_t op = 0;
    refill(b, lane);
    {   // zlib header: CM = 8, no preset dictionary, header check
        const uint32_t cmf = take(b, 8), flg = take(b, 8);
        if ((cmf & 15u) != 8u || ((cmf << 8) | flg) % 31u != 0u || (flg & 32u)) st = 1;
    }
    bool last = false;
    while (!st && !last) {
        refill(b, lane);
        last = take(b, 1) != 0;
        const uint32_t type = take(b, 2);
        if (type == 0) {                                   // stored
            take(b, b.cnt & 7);
            refill(b, lane);
            const uint32_t len = take(b, 16), nlen = take(b, 16);
            if ((len ^ 0xFFFFu) != nlen) { st = 2; break; }
            // the bit buffer holds whole bytes now: continue at the byte position
            const uint32_t start = bits_bytepos(b);
            if (start + len > J.zlen || op + len > out_len) { st = 2; break; }
            for (uint32_t i = lane; i < len; i += 32) out[op + i] = __ldg(J.z + start + i);
            op += len;
            const uint32_t np = start + len;
            bits_init(b, J.z, J.zlen, np, lane);
            refill(b, lane);
            take(b, 8 * (int)(np & 3u));                  // bytes of the first word that belong to the stored data
            __syncwarp();
            continue;
        }
        if (type == 3) { st = 2; break; }
        if (type == 1) {                                   // fixed codes
            for (int i = lane; i < 288; i += 32) T.len[i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
            __syncwarp();
            if (!build_tables(T.len, 288, T.lit, TB_LIT, T.lit_cnt, T.lit_sym, lane)) { st = 3; break; }
            __syncwarp();
            for (int i = lane; i < 30; i += 32) T.len[i] = 5;
            __syncwarp();
            if (!build_tables(T.len, 30, T.dist, TB_DIST, T.dist_cnt, T.dist_sym, lane)) { st = 3; break; }
        } else {                                           // dynamic codes
            const int hlit = (int)take(b, 5) + 257, hdist = (int)take(b, 5) + 1, hclen = (int)take(b, 4) + 4;
            if (hlit > 286 || hdist > 30) { st = 2; break; }
            for (int i = lane; i < 19; i += 32) T.len[i] = 0;
            __syncwarp();
            for (int i = 0; i < hclen; i++) {
                refill(b, lane);
                const uint32_t v = take(b, 3);
                if (lane == 0) T.len[c_clen_order[i]] = (uint8_t)v;
            }
            __syncwarp();
            // the code-length code reuses the distance tables as scratch
            if (!build_tables(T.len, 19, T.dist, 7, T.dist_cnt, T.dist_sym, lane)) { st = 3; break; }
            __syncwarp();
            int i = 0, prev = 0;
            while (i < hlit + hdist) {
                refill(b, lane);
                const int s = decode_sym(b, T.dist, 7, T.dist_cnt, T.dist_sym);
                if (s < 0) { st = 3; break; }
                if (s < 16) {
                    if (lane == 0) T.len[i] = (uint8_t)s;
                    prev = s;
                    i++;
                } else {
                    int rep, val = 0;
                    if (s == 16) { if (i == 0) { st = 3; break; } val = prev; rep = 3 + (int)take(b, 2); }
                    else if (s == 17) rep = 3 + (int)take(b, 3);
                    else rep = 11 + (int)take(b, 7);
                    if (i + rep > hlit + hdist) { st = 3; break; }
                    if (lane == 0)
                        for (int k = 0; k < rep; k++) T.len[i + k] = (uint8_t)val;
                    prev = val;
                    i += rep;
                }
            }
            if (st) break;
            __syncwarp();
            // distance lengths are copied aside before the literal table is built over T.len[0..hlit)
            uint8_t dl = 0;
            if (lane < hdist) dl = T.len[hlit + lane];
            __syncwarp();
            if (T.len[256] == 0) { st = 3; break; }
            if (!build_tables(T.len, hlit, T.lit, TB_LIT, T.lit_cnt, T.lit_sym, lane)) { st = 3; break; }
            __syncwarp();
            if (lane < 32) T.len[lane] = lane < hdist ? dl : 0;
            __syncwarp();
            if (!build_tables(T.len, hdist, T.dist, TB_DIST, T.dist_cnt, T.dist_sym, lane)) { st = 3; break; }
        }
        __syncwarp();
        // ---- symbols of the block ----
        for (;;) {
            refill(b, lane);
#ifndef ABUB_K9_SCALAR
            {   // ---- literal run.  Lane j decodes the codes that WOULD start at bit j of the window (two of them, so that it knows
                // its 2-hop successor without talking to anyone); the codes that really start there are the chain
                // 0 -> 0+len(0) -> ..., marked by pointer doubling over the warp (one REDUX per round, the SHFL that squares the
                // successor map runs beside it).  Every literal on the chain is stored by its own lane; the chain stops at
                // the first length code, end-of-block, long (> TB_LIT bits) code or at the end of the window, where the
                // one-symbol path takes over.
                const uint32_t x = __funnelshift_r((uint32_t)b.buf, (uint32_t)(b.buf >> 32), lane);
                const uint32_t e = T.lit[x & ((1u << TB_LIT) - 1u)];
                const int l = (int)(e & 15u);
                const bool lit = l != 0 && e < (256u << 4) && lane + l <= b.cnt;
                const int cons = lit ? lane + l : lane;             // bits consumed if the chain ends at this lane
                const int s1 = (lit && cons < 32) ? cons : 32;       // successor; 32: none inside the window
                const uint32_t e2 = T.lit[(uint32_t)(b.buf >> cons) & ((1u << TB_LIT) - 1u)];
                const int p2 = s1 + (int)(e2 & 15u);
                int succ = (s1 < 32 && (e2 & 15u) != 0u && e2 < (256u << 4) && p2 <= b.cnt && p2 < 32) ? p2 : 32;   // 2 hops
                const uint32_t term = __ballot_sync(0xffffffffu, s1 >= 32);          // where a chain would end
                const int h1 = __shfl_sync(0xffffffffu, s1, 0);
                uint32_t M = 1u | (h1 < 32 ? 1u << h1 : 0u);
                while (!(M & term)) {                                // the chain from 0 ends at exactly one such lane
                    const int t = __shfl_sync(0xffffffffu, succ, succ & 31);
                    M |= __reduce_or_sync(0xffffffffu, (((M >> lane) & 1u) && succ < 32) ? (1u << succ) : 0u);
                    succ = succ < 32 ? t : 32;
                }
                const uint32_t Mlit = M & __ballot_sync(0xffffffffu, lit);
                const int last_lane = 31 - __clz((int)M);
                const int used = __shfl_sync(0xffffffffu, cons, last_lane);
                const uint32_t n = (uint32_t)__popc(Mlit);
                if (op + n > out_len) { st = 4; break; }
                if ((Mlit >> lane) & 1u) out[op + (uint32_t)__popc(Mlit & ((1u << lane) - 1u))] = (uint8_t)(e >> 4);
                op += n;
                b.buf >>= used;
                b.cnt -= used;
                if (used > last_lane) continue;                      // ran off the window on a literal: next window
                refill(b, lane);
            }
#endif
            const int s = decode_sym(b, T.lit, TB_LIT, T.lit_cnt, T.lit_sym);
            if (s < 256) {
                if (s < 0) { st = 3; break; }
                if (op >= out_len) { st = 4; break; }
                if (lane == 0) out[op] = (uint8_t)s;
                op++;
                continue;
            }
            if (s == 256) break;
            if (s > 285) { st = 3; break; }
            const uint32_t len = c_len_base[s - 257] + take(b, c_len_extra[s - 257]);
            refill(b, lane);                                // 33 bits: distance code (<= 15) + extra bits (<= 13)
            const int ds = decode_sym(b, T.dist, TB_DIST, T.dist_cnt, T.dist_sym);
            if (ds < 0 || ds > 29) { st = 3; break; }
            const uint32_t dist = c_dist_base[ds] + take(b, c_dist_extra[ds]);
            if (dist > op || op + len > out_len) { st = 4; break; }
            __syncwarp();                                   // literals stored by lane 0 become visible to the copy
            const uint8_t* src = out + op - dist;
            for (uint32_t i = lane; i < len; i += 32) out[op + i] = src[i < dist ? i : i % dist];
            __syncwarp();
            op += len;
        }
        if (b.overrun) st = 5;
    }
    if (!st && op != out_len) st = 4;
    if (lane == 0) status[job] = st;
}

__device__ __forceinline__ int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// One warp per frame; shared memory per warp: the filtered row and the reconstructed previous row.
__global__ void __launch_bounds__(32 * K9_WARPS) k9_unfilter_kernel(const PngDev* __restrict__ jobs, int njobs, int rows, int cols,
                                                                    int* __restrict__ status) {
    extern __shared__ uint8_t sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int job = blockIdx.x * K9_WARPS + warp;
    if (job >= njobs) return;
    const PngDev J = jobs[job];
    const int pitch = (cols + 15) & ~15;
    uint8_t* cur = sm + (size_t)warp * 2 * pitch;
    uint8_t* prev = cur + pitch;
    if (status[job] != 0) {                                 // undecodable frame: zero-filled, the host is told
        for (size_t i = lane; i < (size_t)rows * cols; i += 32) J.dst[i] = 0;
        return;
    }
    for (int i = lane; i < cols; i += 32) prev[i] = 0;
    int bad = 0;
    for (int y = 0; y < rows; y++) {
        const uint8_t* line = J.filt + (size_t)y * (cols + 1);
        const int ft = line[0];
        __syncwarp();
        for (int i = lane; i < cols; i += 32) cur[i] = line[1 + i];
        __syncwarp();
        if (ft == 2) {
            for (int i = lane; i < cols; i += 32) cur[i] = (uint8_t)(cur[i] + prev[i]);
        } else if (ft == 1) {
            // prefix sum modulo 256 in chunks of 32 pixels
            uint32_t carry = 0;
            for (int i0 = 0; i0 < cols; i0 += 32) {
                const int i = i0 + lane;
                uint32_t v = i < cols ? cur[i] : 0;
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
                    if (lane >= o) v += t;
                }
                v += carry;
                if (i < cols) cur[i] = (uint8_t)v;
                carry = __shfl_sync(0xffffffffu, v, 31) & 255u;
            }
        } else if (ft == 3 || ft == 4) {
            if (lane == 0) {
                int a = 0, c = 0;
                if (ft == 3) {
                    for (int i = 0; i < cols; i++) {
                        a = (cur[i] + ((a + prev[i]) >> 1)) & 255;
                        cur[i] = (uint8_t)a;
                    }
                } else {
                    for (int i = 0; i < cols; i++) {
                        const int bb = prev[i];
                        a = (cur[i] + paeth(a, bb, c)) & 255;
                        c = bb;
                        cur[i] = (uint8_t)a;
                    }
                }
            }
        } else if (ft != 0) {
            bad = 1;
        }
        __syncwarp();
        uint8_t* dst = J.dst + (size_t)y * cols;
        for (int i = lane; i < cols; i += 32) dst[i] = cur[i];
        uint8_t* t = cur; cur = prev; prev = t;
    }
    if (bad && lane == 0) status[job] = 6;
}

}  // namespace

// ------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------
struct PngArena {
    PngDev* d_jobs = nullptr;
    int* d_status = nullptr;
    uint8_t* d_filt = nullptr;
    int cap = 0;
    std::vector<PngDev> h_jobs;
    std::vector<int> h_status;
};

void abub_png_free(abub_ctx* ctx) {
    PngArena* a = reinterpret_cast<PngArena*>(ctx->png);
    if (!a) return;
    cudaFree(a->d_jobs); cudaFree(a->d_status); cudaFree(a->d_filt);
    delete a;
    ctx->png = nullptr;
}

extern "C" int abub_event_decode_png(abub_ctx* ctx, const abub_png_job* jobs, int n, int* ok) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, jobs && n >= 0 && ok, "abub_event_decode_png");
    if (n == 0) return ABUB_OK;
    cudaSetDevice(ctx->device);
    if (!ctx->png) ctx->png = new PngArena();
    PngArena* a = reinterpret_cast<PngArena*>(ctx->png);
    const size_t flen = (size_t)ctx->rows * (ctx->cols + 1);
    for (int i = 0; i < n; i++) {
        const abub_png_job& j = jobs[i];
        ABUB_REQUIRE(ctx, j.slot >= 0 && j.slot < ctx->max_slots && j.cam >= 0 && j.cam < ctx->ncams && j.frame >= 0 &&
                              j.frame < ctx->max_frames, "png job slot/cam/frame");
        ABUB_REQUIRE(ctx, j.zlen >= 2 && (size_t)j.zlen <= ctx->npix, "png job: stream longer than a frame");
        ABUB_REQUIRE(ctx, ctx->npix % 4 == 0, "on-device PNG decode needs rows*cols to be a multiple of 4 (word-aligned frame regions)");
        ABUB_REQUIRE(ctx, ctx->slots[j.slot].h_pinned != nullptr, "slot has no pinned buffer (call abub_pinned_acquire)");
    }
    if (n > a->cap) {
        cudaFree(a->d_jobs); cudaFree(a->d_status); cudaFree(a->d_filt);
        a->d_jobs = nullptr; a->d_status = nullptr; a->d_filt = nullptr; a->cap = 0;
        const int cap = n + n / 4 + 16;
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&a->d_jobs, (size_t)cap * sizeof(PngDev)));
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&a->d_status, (size_t)cap * sizeof(int)));
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&a->d_filt, (size_t)cap * flen));
        a->cap = cap;
    }
    cudaStream_t st = ctx->stream;
    a->h_jobs.resize(n);
    a->h_status.assign(n, -1);
    for (int i = 0; i < n; i++) {
        const abub_png_job& j = jobs[i];
        abub_slot_state& s = ctx->slots[j.slot];
        const size_t off = ((size_t)j.cam * ctx->max_frames + j.frame) * ctx->npix;
        // the compressed stream waits at the start of the frame's own staging region and is sent to the start of the
        // frame's device memory; the inflated scanlines go to the scratch, the pixels then overwrite the stream
        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(s.d_frames + off, s.h_pinned + off, j.zlen, cudaMemcpyHostToDevice, st));
        a->h_jobs[i] = PngDev{s.d_frames + off, j.zlen, a->d_filt + (size_t)i * flen, s.d_frames + off};
        s.hist_valid = false;
    }
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(a->d_jobs, a->h_jobs.data(), (size_t)n * sizeof(PngDev), cudaMemcpyHostToDevice, st));
    const int nb = (n + K9_WARPS - 1) / K9_WARPS;
    ABUB_L(ctx), k9_inflate_kernel<<<nb, 32 * K9_WARPS, 0, st>>>(a->d_jobs, n, (uint32_t)flen, a->d_status);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    const size_t smem = (size_t)K9_WARPS * 2 * ((ctx->cols + 15) & ~15);
    ABUB_REQUIRE(ctx, smem <= 200 * 1024, "frame too wide for the on-device PNG unfilter");
    ABUB_CHECK_CUDA(ctx, cudaFuncSetAttribute(k9_unfilter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ABUB_L(ctx), k9_unfilter_kernel<<<nb, 32 * K9_WARPS, smem, st>>>(a->d_jobs, n, ctx->rows, ctx->cols, a->d_status);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(a->h_status.data(), a->d_status, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, st));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(st));
    for (int i = 0; i < n; i++) ok[i] = a->h_status[i] == 0;
    return ABUB_OK;
}


// ---- bottom-up BMP pixel arrays: rows turned over in place ---------------------------------------------------------
namespace {
constexpr int FLIP_REFS = 256;                     // frame references per launch: they travel as a kernel argument (3 KB),
struct FlipRefs { abub_frame_ref r[FLIP_REFS]; };  // so the call needs no device buffer, no copy and no synchronisation

// grid: x = chunks of a frame's upper half, y = frame.  One thread swaps 4 bytes of row r with 4 bytes of row rows-1-r.
__global__ void __launch_bounds__(256) k9_flip_rows_kernel(uint8_t* const* __restrict__ slot_ptrs, const __grid_constant__ FlipRefs refs,
                                                           int rows, int cols, int max_frames) {
    const abub_frame_ref f = refs.r[blockIdx.y];
    uint32_t* const img = reinterpret_cast<uint32_t*>(slot_ptrs[f.slot] + ((size_t)f.cam * max_frames + f.frame) * ((size_t)rows * cols));
    const int wpr = cols >> 2;                                   // words per row
    const int half = rows >> 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < half * wpr; i += gridDim.x * blockDim.x) {
        const int r = i / wpr, w = i - r * wpr;
        uint32_t* const a = img + (size_t)r * wpr + w;
        uint32_t* const b = img + (size_t)(rows - 1 - r) * wpr + w;
        const uint32_t va = *a, vb = *b;
        *a = vb;
        *b = va;
    }
}
}  // namespace

extern "C" int abub_event_flip_rows(abub_ctx* ctx, const abub_frame_ref* frames, int n) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, n >= 0 && (frames || n == 0), "abub_event_flip_rows");
    ABUB_REQUIRE(ctx, (ctx->cols & 3) == 0, "abub_event_flip_rows needs cols % 4 == 0");
    for (int i = 0; i < n; i++)
        ABUB_REQUIRE(ctx, frames[i].slot >= 0 && frames[i].slot < ctx->max_slots && frames[i].cam >= 0 && frames[i].cam < ctx->ncams &&
                              frames[i].frame >= 0 && frames[i].frame < ctx->max_frames, "frame reference out of range");
    cudaSetDevice(ctx->device);
    FlipRefs refs;
    for (int i0 = 0; i0 < n; i0 += FLIP_REFS) {
        const int cnt = n - i0 < FLIP_REFS ? n - i0 : FLIP_REFS;
        memcpy(refs.r, frames + i0, (size_t)cnt * sizeof(abub_frame_ref));
        ABUB_L(ctx), k9_flip_rows_kernel<<<dim3(32, cnt), 256, 0, ctx->stream>>>(ctx->d_slot_ptrs, refs, ctx->rows, ctx->cols, ctx->max_frames);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    return ABUB_OK;
}
