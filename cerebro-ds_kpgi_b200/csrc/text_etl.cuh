// text_etl.cuh — gpfdist TEXT unload -> float32 arrays on the GPU.
//
// Replaces the parser of the reference's filesystem ETL
// (/root/reference/cerebro_gpdb/etl_imagenet.py:30-37, etl_criteo.py:28-37) for files written by
// /root/reference/cerebro_gpdb/unload_imagenet.sql:1-14 (FORMAT 'TEXT', DELIMITER '|'): one row per
// line, `id|{v,v,...}|label`, the feature column a (possibly nested) brace array of decimal numbers.
//   reference:  literal_eval(text with braces -> brackets) -> Python numbers -> np.asarray(float32)
//   here:       every numeric token is parsed by one thread; decimal -> double is exact
//               (Eisel-Lemire: 64-bit mantissa times a 128-bit power of five, correctly rounded; Clinger's
//               one-rounding multiply/divide as the back-up), then double -> float32 (round to nearest
//               even), i.e. the same double rounding as float(token) followed by np.float32.
// Refused with CDS_ERR_TEXT_NUMBER (there is no slow path on the CPU to fall back to): more than 19
// significant digits, decimal exponents outside [-80, 60] after the digits are shifted into the mantissa,
// NaN, Infinity and malformed literals (literal_eval raises on the last three as well).
//
// Rows are uniform: every line holds `per_row` numeric tokens (id, features..., label) in that order,
// which is checked at every line end.
#pragma once
#include "cds_common.cuh"

namespace cds {
namespace {   // internal linkage: with relocatable device code (-rdc) the optimiser then treats the file as a whole program

constexpr uint32_t kTextTile = 4096;          // characters per CTA tile
constexpr uint32_t kTextThreads = 256;        // 16 characters per thread
constexpr uint32_t kTextHalo = 64;            // characters of the next tile staged with this one

__device__ __forceinline__ bool is_num_char(uint32_t c)
{
    return (c >= '0' && c <= '9') || c == '-' || c == '+' || c == '.' || (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z');
}

struct TextState {
    unsigned long long err_inv;   // ~(smallest error key): position of the first bad token / line
    unsigned long long n_tokens;
    unsigned long long n_lines;
    unsigned long long pad;
};

// The 16 characters of this thread (zero past the end) and the character before them.
__device__ __forceinline__ void load16(const uint8_t *__restrict__ text, unsigned long long base, unsigned long long n,
                                       uint32_t w[4], uint32_t &prev)
{
    if (base + 16 <= n && ((reinterpret_cast<uintptr_t>(text) + base) & 15u) == 0) {
        const uint4 q = __ldg(reinterpret_cast<const uint4 *>(text + base));
        w[0] = q.x; w[1] = q.y; w[2] = q.z; w[3] = q.w;
    } else {
        w[0] = w[1] = w[2] = w[3] = 0;
        for (uint32_t k = 0; k < 16 && base + k < n; k++) w[k >> 2] |= (uint32_t)text[base + k] << ((k & 3) * 8);
    }
    // the previous thread holds my predecessor character; the first lane of a warp reads it
    const uint32_t up = __shfl_up_sync(0xffffffffu, w[3] >> 24, 1);
    prev = (threadIdx.x & 31) ? up : ((base > 0 && base <= n) ? (uint32_t)text[base - 1] : (uint32_t)'\n');
}

// Pass 1: per tile, how many tokens start and how many lines end inside it.
__global__ void __launch_bounds__(kTextThreads)
k_text_count(const uint8_t *__restrict__ text, unsigned long long n, uint32_t *__restrict__ tile_tokens,
             uint32_t *__restrict__ tile_lines)
{
    __shared__ uint32_t sh[2][kTextThreads / 32];
    const unsigned long long base = (unsigned long long)blockIdx.x * kTextTile + threadIdx.x * 16ull;
    uint32_t tok = 0, nl = 0, w[4], prev;
    load16(text, base, n, w, prev);
#pragma unroll
    for (uint32_t k = 0; k < 16; k++) {
        const uint32_t c = (w[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
        const bool in = base + k < n;
        tok += (in && is_num_char(c) && !is_num_char(prev)) ? 1u : 0u;
        nl += (in && c == '\n') ? 1u : 0u;
        prev = c;
    }
    for (int d = 16; d > 0; d >>= 1) { tok += __shfl_xor_sync(0xffffffffu, tok, d); nl += __shfl_xor_sync(0xffffffffu, nl, d); }
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = tok; sh[1][threadIdx.x >> 5] = nl; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t a = 0, b = 0;
        for (uint32_t w = 0; w < kTextThreads / 32; w++) { a += sh[0][w]; b += sh[1][w]; }
        tile_tokens[blockIdx.x] = a;
        tile_lines[blockIdx.x] = b;
    }
}

// Pass 2: exclusive scans of the two per-tile counts (blockIdx.y selects which).
__global__ void __launch_bounds__(kScanThreads)
k_text_scan(uint32_t *tile_tokens, uint32_t *tile_lines, uint32_t n_tiles, ScanWs ws0, ScanWs ws1, TextState *st,
            unsigned long long n, const uint8_t *__restrict__ text)
{
    bool last;
    if (blockIdx.y == 0) {
        uint32_t *x = tile_tokens;
        const uint32_t t = device_exclusive_scan<uint32_t, 4>([&](uint32_t i) { return x[i]; }, x, n_tiles, ws0, true, &last);
        if (last && threadIdx.x == 0) st->n_tokens = t;
    } else {
        uint32_t *x = tile_lines;
        const uint32_t t = device_exclusive_scan<uint32_t, 4>([&](uint32_t i) { return x[i]; }, x, n_tiles, ws1, true, &last);
        // a last line without a trailing newline still counts (readlines() returns it)
        if (last && threadIdx.x == 0) st->n_lines = t + ((n > 0 && text[n - 1] != '\n') ? 1u : 0u);
    }
}

#include "pow5_table.inc"

// Eisel-Lemire: w * 10^q -> the correctly rounded binary64, from one (rarely two) 64x64->128-bit products with a
// 128-bit approximation of 5^q.  Returns false in the few cases the method cannot decide (the caller then tries
// Clinger's exact path or refuses).  Checked against Python's float() on 700 k random and near-halfway inputs
// by the generator-side model of this routine before it was written here.
__device__ inline bool eisel_lemire(unsigned long long w, int q, double *out)
{
    if (w == 0) { *out = 0.0; return true; }
    if (q < CDS_POW5_QMIN || q > CDS_POW5_QMAX) return false;
    const int lz = __clzll((long long)w);
    w <<= lz;
    const unsigned long long hi5 = kPow5[q - CDS_POW5_QMIN][0], lo5 = kPow5[q - CDS_POW5_QMIN][1];
    unsigned long long fh = __umul64hi(w, hi5), fl = w * hi5;
    if ((fh & 0x1FFull) == 0x1FFull) {                       // the 55 bits we need are not settled yet: refine
        const unsigned long long sh = __umul64hi(w, lo5);
        fl += sh;
        if (sh > fl) fh++;
    }
    if (fl == ~0ull && !(q >= -27 && q <= 55)) return false;
    const int upperbit = (int)(fh >> 63), s = upperbit + 9;  // 64 - 52 - 3
    unsigned long long mant = fh >> s;
    int power2 = (((152170 + 65536) * q) >> 16) + 63 + upperbit - lz + 1023;
    if (power2 <= 0) return false;                           // binary64 subnormal: outside what the unloads hold
    if (fl <= 1ull && q >= -4 && q <= 23 && (mant & 3ull) == 1ull && (mant << s) == fh) mant &= ~1ull;   // exact tie: to even
    mant += mant & 1ull;
    mant >>= 1;
    if (mant >= (2ull << 52)) { mant = 1ull << 52; power2++; }
    mant &= ~(1ull << 52);
    if (power2 >= 0x7FF) return false;
    *out = __longlong_as_double((long long)(((unsigned long long)power2 << 52) | mant));
    return true;
}

// 10^0 .. 10^22 are exactly representable in binary64
__constant__ double kPow10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15,
                                  1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
__device__ __forceinline__ double pow10_exact(uint32_t k) { return kPow10[k]; }

// Parse the numeric token starting at text[i].  Returns false when it is not a plain decimal literal or
// cannot be converted exactly by the fast path.  *out = float(token) rounded to float32; *as_int = the
// integer value when the token is an integer literal (labels, ids).
__device__ inline bool parse_number(const uint8_t *__restrict__ text, unsigned long long i, unsigned long long n,
                                    float *out, long long *as_int, uint32_t *ndigits_int)
{
    bool neg = false;
    if (i < n && (text[i] == '-' || text[i] == '+')) { neg = text[i] == '-'; i++; }
    unsigned long long m = 0;
    int e10 = 0, sig = 0, digits = 0;
    bool any = false, frac = false, is_int = true;
    for (; i < n; i++) {
        const uint32_t c = text[i];
        if (c >= '0' && c <= '9') {
            any = true;
            digits++;
            if (sig < 19) {
                m = m * 10ull + (c - '0');
                if (m) sig++;
                if (frac) e10--;
            } else {
                if (c != '0') return false;             // a 20th significant digit: not exactly handled
                if (!frac) e10++;
            }
        } else if (c == '.' && !frac) {
            frac = true;
            is_int = false;
        } else {
            break;
        }
    }
    if (!any) return false;
    if (i < n && (text[i] == 'e' || text[i] == 'E')) {
        is_int = false;
        i++;
        bool eneg = false;
        if (i < n && (text[i] == '-' || text[i] == '+')) { eneg = text[i] == '-'; i++; }
        int ex = 0, nd = 0;
        for (; i < n && text[i] >= '0' && text[i] <= '9'; i++, nd++) ex = ex < 100000 ? ex * 10 + (text[i] - '0') : ex;
        if (nd == 0) return false;
        e10 += eneg ? -ex : ex;
    }
    if (i < n && is_num_char(text[i])) return false;   // trailing letters / second sign: NaN, Infinity, 1.2.3 ...
    double d;
    if (m == 0) {
        d = 0.0;
    } else if (eisel_lemire(m, e10, &d)) {
        // correctly rounded binary64 of m * 10^e10
    } else if (m < (1ull << 53) && e10 >= 0 && e10 <= 22) {
        d = (double)m * pow10_exact((uint32_t)e10);          // Clinger's exact path: one IEEE rounding
    } else if (m < (1ull << 53) && e10 < 0 && e10 >= -22) {
        d = (double)m / pow10_exact((uint32_t)-e10);
    } else {
        return false;
    }
    if (is_int && m == 0) neg = false;                   // "-0" is the Python int 0: np.float32(0) is +0.0 ("-0.0" stays -0.0)
    *out = (float)(neg ? -d : d);                        // double -> float32, round to nearest even
    *as_int = is_int ? (neg ? -(long long)m : (long long)m) : 0;
    *ndigits_int = is_int ? (uint32_t)digits : 0u;
    return true;
}

// Pass 3: every token start is parsed by the thread that owns its character.
//   token t of the file: row = t / per_row, k = t % per_row; k == 0 -> id, k == per_row-1 -> label,
//   otherwise feature k-1.  label_digits != 0: the label is a string of that many digits, each an int
//   (etl_criteo.py:35-36).  Every line end must fall on a row boundary of the token stream.
__global__ void __launch_bounds__(kTextThreads)
k_text_parse(const uint8_t *__restrict__ text, unsigned long long n, const uint32_t *__restrict__ tile_tokens,
             const uint32_t *__restrict__ tile_lines, uint32_t per_row, uint32_t label_digits,
             float *__restrict__ features, long long *__restrict__ labels, long long *__restrict__ ids,
             unsigned long long cap_rows, TextState *st)
{
    __shared__ uint32_t sh_tok[kTextThreads / 32], sh_nl[kTextThreads / 32];
    __shared__ __align__(16) uint8_t tile[kTextTile + kTextHalo];           // this tile's characters + the start of the next
    const unsigned long long tile_base = (unsigned long long)blockIdx.x * kTextTile;
    const unsigned long long base = tile_base + threadIdx.x * 16ull;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t flags = 0, nls = 0, w[4], prev;                                 // bit k: token start / newline at base+k
    load16(text, base, n, w, prev);
    *reinterpret_cast<uint4 *>(tile + threadIdx.x * 16u) = make_uint4(w[0], w[1], w[2], w[3]);
    if (threadIdx.x < kTextHalo) {
        const unsigned long long h = tile_base + kTextTile + threadIdx.x;
        tile[kTextTile + threadIdx.x] = h < n ? text[h] : (uint8_t)'\n';
    }
#pragma unroll
    for (uint32_t k = 0; k < 16; k++) {
        const uint32_t c = (w[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
        const bool in = base + k < n;
        if (in && is_num_char(c) && !is_num_char(prev)) flags |= 1u << k;
        if (in && c == '\n') nls |= 1u << k;
        prev = c;
    }
    // ranks inside the tile
    uint32_t tcnt = __popc(flags), ncnt = __popc(nls), tinc = tcnt, ninc = ncnt;
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, tinc, d), b = __shfl_up_sync(0xffffffffu, ninc, d);
        if (lane >= (uint32_t)d) { tinc += a; ninc += b; }
    }
    if (lane == 31) { sh_tok[warp] = tinc; sh_nl[warp] = ninc; }
    __syncthreads();
    uint32_t toff = 0, noff = 0;
    for (uint32_t w = 0; w < warp; w++) { toff += sh_tok[w]; noff += sh_nl[w]; }
    unsigned long long t = (unsigned long long)tile_tokens[blockIdx.x] + toff + tinc - tcnt;   // index of my first token
    unsigned long long l = (unsigned long long)tile_lines[blockIdx.x] + noff + ninc - ncnt;    // lines ended before my chars
    const uint32_t nfeat = per_row - 2u;
    const unsigned long long staged = min((unsigned long long)(kTextTile + kTextHalo), n - min(n, tile_base));
    // Every lane walks its own token starts (lowest first), so that the j-th tokens of all lanes are parsed together.
    uint32_t todo = flags;
    while (todo) {
        const uint32_t k = (uint32_t)__ffs((int)todo) - 1u;
        todo &= todo - 1u;
        const unsigned long long i = base + k;
        const unsigned long long row = t / per_row;
        const uint32_t col = (uint32_t)(t - row * per_row);
        float v;
        long long iv;
        uint32_t nd;
        // tokens are parsed out of shared memory; one that could run past the staged halo goes to global memory
        const uint32_t li = threadIdx.x * 16u + k;
        const bool ok = (li + 48u <= staged || staged == n - tile_base) ? parse_number(tile, li, staged, &v, &iv, &nd)
                                                                       : parse_number(text, i, n, &v, &iv, &nd);
        if (!ok) {
            atomicMax(&st->err_inv, ~make_err_key(0, i >> 4, 0, (uint32_t)(i & 15), CDS_ERR_TEXT_NUMBER));
        } else if (row < cap_rows) {
            if (col == 0) { if (ids) ids[row] = iv; }
            else if (col <= nfeat) features[row * nfeat + (col - 1u)] = v;
            else if (!label_digits) labels[row] = iv;
            else {
                if (nd != label_digits) atomicMax(&st->err_inv, ~make_err_key(1, i >> 4, 0, (uint32_t)(i & 15), CDS_ERR_SHAPE));
                else {
                    unsigned long long x = (unsigned long long)iv;
                    for (int q = (int)label_digits - 1; q >= 0; q--) { labels[row * label_digits + q] = (long long)(x % 10ull); x /= 10ull; }
                }
            }
        } else {
            atomicMax(&st->err_inv, ~make_err_key(2, 0, 0, 0, CDS_ERR_CAPACITY));
        }
        t++;
    }
    // line ends: the tokens seen before a newline must be exactly (lines ended + 1) rows worth
    const unsigned long long t0 = t - tcnt;
    todo = nls;
    while (todo) {
        const uint32_t k = (uint32_t)__ffs((int)todo) - 1u;
        todo &= todo - 1u;
        const unsigned long long i = base + k;
        const unsigned long long t_at = t0 + __popc(flags & ((1u << k) - 1u));
        const unsigned long long l_at = l + __popc(nls & ((1u << k) - 1u));
        if (t_at != (l_at + 1ull) * per_row)
            atomicMax(&st->err_inv, ~make_err_key(1, i >> 4, 0, (uint32_t)(i & 15), CDS_ERR_SHAPE));
    }
}

}  // namespace
}  // namespace cds
