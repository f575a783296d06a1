// cds_da.cu — C ABI of libcds_da.so (include/cds_da.h): handle management and kernel sequencing.
//
// The decode of one segment is a fixed sequence of launches on the caller's stream with no host
// round trip in between: page index -> row parse -> value index -> TOAST walk -> plan (ranks and
// prefix sums) -> resolve (checks + copy descriptors) -> gather -> pglz.  Counts stay on the
// device; cds_finish() reads them back once.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>

#include "cast_pack.cuh"
#include "cds_common.cuh"
#include "page_scan.cuh"
#include "pglz_decode.cuh"
#include "pglz_wide.cuh"
#include "stream_decode.cuh"
#include "text_etl.cuh"
#include "toast_gather.cuh"
#include "util_kernels.cuh"

using namespace cds;

static thread_local char g_last_error[512] = "";
static unsigned long long g_launches = 0;   // kernels launched by this library (bench.py reports it)
#define LAUNCHED(n) (g_launches += (n))
static uint32_t g_wide_threshold = 1u << 20;   // cds_set_wide_pglz_threshold
static uint64_t g_wide_store_limit = ~0ull;   // cds_set_wide_store_limit: most bytes a handle's store of resolved blocks may take (0: none)
static int g_wide_force = 0;                   // ... its second argument: skip the cost model
static unsigned long long g_wide_decoded = 0;  // values decoded by the wide decoder so far (cds_wide_decoded_count)
static unsigned long long g_streamed = 0, g_stream_redone = 0;   // cds_decode_host calls that took the streamed path / were decoded again

static int cuda_fail(cudaError_t e, const char *what)
{
    snprintf(g_last_error, sizeof(g_last_error), "%s: %s", what, cudaGetErrorString(e));
    return CDS_ERR_CUDA;
}
#define CU(call)                                              \
    do {                                                      \
        cudaError_t e__ = (call);                             \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
    } while (0)

// Every entry point that makes `device` current restores the caller's device when it returns.
struct DeviceScope {
    int prev = -1;
    cudaError_t err = cudaSuccess;
    explicit DeviceScope(int device)
    {
        int cur = -1;
        err = cudaGetDevice(&cur);
        if (err == cudaSuccess && cur != device) {
            err = cudaSetDevice(device);
            if (err == cudaSuccess) prev = cur;
        }
    }
    ~DeviceScope() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceScope(const DeviceScope &) = delete;
    DeviceScope &operator=(const DeviceScope &) = delete;
};
#define ON_DEVICE(dev)                                                         \
    DeviceScope device_scope__(dev);                                           \
    if (device_scope__.err != cudaSuccess) return cuda_fail(device_scope__.err, "cudaSetDevice")

// Multiprocessor count of `device` (queried once per device).
static int sm_count(int device)
{
    static int cached[64] = {0};
    if (device < 0 || device >= 64) return 148;
    if (cached[device] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) { cudaGetLastError(); n = 148; }
        cached[device] = n;
    }
    return cached[device];
}
static int current_sm_count()
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return 148; }
    return sm_count(dev);
}

constexpr uint32_t kStreamDecoders = 4;
constexpr uint32_t kTimedStages = 4;   // 0 gather, 1 pglz stage (classify + wide + warp decoder), 2 / 3 the wide block passes

struct cds_segment {
    int device = 0;
    int num_sms = 148;
    uint64_t cap_heap_pages = 0, cap_toast_pages = 0, cap_rows = 0, cap_chunks = 0, cap_scratch = 0;
    uint32_t hmask = 0;
    // inputs of the current decode
    const uint8_t *d_heap = nullptr, *d_toast = nullptr;
    uint64_t n_heap_pages = 0, n_toast_pages = 0;
    bool have_rows = false, have_chunks = false, have_plan = false, have_ops = false;
    int coop_blocks = 0;            // co-resident grid for k_plan_fused (0: cooperative launch unavailable)
    int ind_format = 0;             // cds_set_output_format: 0 stored bytes, 1 f32 -> bf16, 2 u8 -> f32 through the table
    float *d_lut = nullptr;         // cds_set_normalize_lut: [channels][256]
    int wide_mode = 2;              // cds_set_wide_pglz: 0 never, 1 always and without the cost model, 2 automatic, 3 always
    bool wide_hint = true;          // mode 2: the last summary saw values go wide (or there is none yet) -> enqueue the wide kernels with the decode
    bool wide_launched = false;     // this decode classified its jobs: cds_finish reads the counters (and decodes what is left)
    WideArea last_wa;               // the work area as this decode's classification used it
    uint8_t *wide_mem = nullptr;    // work area of the intra-value parallel pglz decoder (pglz_wide.cuh)
    uint16_t *wide_store = nullptr; // its store of resolved blocks (2 bytes per decoded byte of the wide values of one decode), sized by cds_finish
    uint64_t wide_store_blks = 0;
    WideArea wide;
    uint32_t lut_channels = 0;
    // device buffers
    uint8_t *stage_heap = nullptr, *stage_toast = nullptr;
    uint32_t *heap_base = nullptr, *toast_base = nullptr, *rows = nullptr;
    unsigned long long *htab = nullptr;
    uint2 *dir_id = nullptr, *dir_loc = nullptr;
    uint32_t *dir_ref = nullptr, *chunk_flag = nullptr;
    uint32_t *vfirst = nullptr, *vcount = nullptr, *vseq0 = nullptr;
    unsigned long long *vbytes = nullptr, *vscr = nullptr;
    uint32_t *inline_flag = nullptr, *row_first = nullptr, *pos_of_row = nullptr, *row_at_pos = nullptr;
    unsigned long long *ind_len = nullptr, *dep_len = nullptr, *pos_ind = nullptr, *pos_dep = nullptr;
    unsigned long long *row_ind_off = nullptr, *row_dep_off = nullptr;
    CopyOp *ops = nullptr;
    PglzJob *jobs = nullptr;
    uint8_t *scratch = nullptr;
    // one allocation cleared by ONE memset per decode: [DevState][scan tickets+status][hash table]
    uint8_t *zero_region = nullptr;
    uint64_t zero_bytes = 0;
    unsigned long long *scan_vals = nullptr;   // aggregates / inclusive prefixes of the scans (not cleared)
    ScanWs ws[7];                               // heap index, TOAST index, inline rank, chunk rank, ind, dep, staging
    DevState *state = nullptr;
    DevState *h_state = nullptr;   // pinned
    // streamed cds_decode_host (stream_decode.cuh): copy stream, one event per page range, the assumed plan
    uint32_t stream_range_pages = 2048;   // cds_set_host_streaming: pages per range (0: off)
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copy[kMaxStreamRanges + 2] = {};
    cudaStream_t dec_stream[kStreamDecoders] = {};           // the warp decoder of each batch runs beside the main stream
    cudaEvent_t ev_batch[kMaxStreamRanges + 1] = {}, ev_dec[kStreamDecoders] = {};
    StreamArgs sargs = {};
    ScanWs ws_range[kMaxStreamRanges];
    uint8_t *stream_mem = nullptr;
    bool stream_pending = false;          // the last decode was streamed and has not been checked by cds_finish yet
    struct { uint64_t n_heap_pages, n_toast_pages; uint32_t isb, dsb; void *out_ind, *out_dep; uint64_t ind_cap, dep_cap; } redo = {};
    // optional CUDA-event timing (cds_timing_enable): kTimedStages event pairs per sample
    cudaEvent_t *ev = nullptr;
    uint32_t ev_cap = 0, ev_used = 0;
    uint8_t *ev_mask = nullptr;     // per sample: which stages were recorded
};

// ---- library -----------------------------------------------------------------------------------

extern "C" int cds_version(void) { return 200; }
extern "C" uint32_t cds_summary_size(void) { return (uint32_t)sizeof(cds_summary); }
static_assert(sizeof(cds_summary) == 80, "cds_summary is part of the ABI: INTEGRATION.md and _lib.py mirror its 80 bytes");

extern "C" const char *cds_status_str(int s)
{
    switch (s) {
    case CDS_OK: return "ok";
    case CDS_ERR_PARTIAL_PAGE: return "FATAL";
    case CDS_ERR_ITEMID_ALIGN: return "Item identifier length do not match";
    case CDS_ERR_TOAST_SPECIAL: return "THERE SHALL NOT BE INDICES";
    case CDS_ERR_TUPLE: return "tuple does not fit its page";
    case CDS_ERR_RAWSIZE: return "raw size does not match!";
    case CDS_ERR_CHUNK_COUNT: return "chunk count does not match va_extsize";
    case CDS_ERR_CHUNK_EXTENDED: return "chunk header is extended";
    case CDS_ERR_MAX_CHUNK: return "max chunk size error";
    case CDS_ERR_CHUNK_SIZE: return "unexpected chunk size";
    case CDS_ERR_FINAL_SIZE: return "final size does not match";
    case CDS_ERR_PGLZ_CORRUPT: return "compressed data is corrupt";
    case CDS_ERR_SHAPE: return "decoded size does not match the sample shape";
    case CDS_ERR_CATALOG_MISS: return "chunk id matches no external pointer";
    case CDS_ERR_MISSING_VALUE: return "a value has no chunks";
    case CDS_ERR_TEXT_NUMBER: return "numeric literal outside the exactly convertible range";
    case CDS_ERR_CUDA: return "CUDA error";
    case CDS_ERR_ARG: return "bad argument";
    case CDS_ERR_CAPACITY: return "capacity exceeded";
    case CDS_ERR_NO_DEVICE: return "no CUDA device";
    default: return "unknown";
    }
}

extern "C" const char *cds_last_error(void) { return g_last_error; }

extern "C" int cds_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" int cds_host_alloc(void **ptr, uint64_t bytes)
{
    if (!ptr) return CDS_ERR_ARG;
    if (cds_device_count() == 0) return CDS_ERR_NO_DEVICE;
    CU(cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocPortable));
    return CDS_OK;
}

extern "C" int cds_host_free(void *ptr)
{
    if (ptr) CU(cudaFreeHost(ptr));
    return CDS_OK;
}

// ---- intra-value parallel pglz (pglz_wide.cuh) -----------------------------------------------------

constexpr uint32_t kPglzCtasPerSm = 2;

// k_pglz_decode needs more dynamic shared memory than the default limit: once per device.
static cudaError_t pglz_prepare(int device)
{
    static bool done[64] = {false};
    if (device >= 0 && device < 64 && done[device]) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(k_pglz_decode, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPglzSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_pglz_decode, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e == cudaSuccess && device >= 0 && device < 64) done[device] = true;
    return e;
}

static cudaError_t wide_prepare()
{
    cudaError_t e = cudaFuncSetAttribute(k_wide_blocks<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWideBlockSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_wide_blocks<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWideBlockSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_wide_blocks<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWideBlockSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_wide_tails<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWideTailsSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_wide_tails<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWideTailsSmem);
    return e;
}

// The 15 kernels behind the classification, on stream `st` (of the four block kernels two return at once: <2> + finish run
// when the decode's blocks fit the store of resolved blocks, <1> + <3> otherwise).
#define CDS_WIDE_PIPELINE(wa, sms, st, BEFORE_B1, AFTER_B1, BEFORE_B3, AFTER_B3)            \
    do {                                                                                    \
        k_wide_maps<<<(sms) * 8u, 256, 0, st>>>(wa);                                        \
        k_wide_super<<<(sms) * 4u, 128, 0, st>>>(wa);                                       \
        k_wide_chain<<<(sms), 32, 0, st>>>(wa);                                             \
        k_wide_entries<<<(sms) * 4u, 128, 0, st>>>(wa);                                     \
        k_wide_outs<<<(sms) * 8u, 256, 0, st>>>(wa);                                        \
        k_wide_scan<<<(sms) * 2u, kScanThreads, 0, st>>>(wa);                               \
        k_wide_check<<<8, 256, 0, st>>>(wa);                                                \
        k_wide_blkmap<<<(sms) * 2u, 256, 0, st>>>(wa);                                      \
        BEFORE_B1;                                                                          \
        k_wide_blocks<2><<<(sms), kWideThreads, kWideBlockSmem, st>>>(wa);                  \
        k_wide_blocks<1><<<(sms), kWideThreads, kWideBlockSmem, st>>>(wa);                  \
        AFTER_B1;                                                                           \
        k_wide_tails<0><<<(sms), 1024, kWideTailsSmem, st>>>(wa);                           \
        k_wide_tails<1><<<(sms), 1024, 0, st>>>(wa);                                        \
        k_wide_tails<2><<<(sms), 1024, kWideTailsSmem, st>>>(wa);                           \
        BEFORE_B3;                                                                          \
        k_wide_blocks<3><<<(sms), kWideThreads, kWideBlockSmem, st>>>(wa);                  \
        k_wide_finish<<<(sms) * 4u, kFinishThreads, 0, st>>>(wa);                           \
        AFTER_B3;                                                                           \
    } while (0)

// Classify the job list (cost model on the device); the jobs taken are blanked, so the warp decoder that runs
// afterwards skips them.  with_pipeline: also decode them (the 15 kernels; `ev4`: optional event pairs around the two
// block passes).  Without it only the classification runs (one small launch): if it does take values,
// kFlagNeedWide is raised and cds_finish runs wide_pipeline() for them.
static int wide_pipeline(WideArea wa, int num_sms, cudaStream_t st, cudaEvent_t *ev4 = nullptr, unsigned ev_flags = cudaEventRecordDefault)
{
    const uint32_t sms = (uint32_t)num_sms;
    CU(cudaMemsetAsync(reinterpret_cast<uint8_t *>(wa.state) + 256, 0, wa.zero_bytes - 256, st));
    CDS_WIDE_PIPELINE(wa, sms, st,
                      if (ev4) cudaEventRecordWithFlags(ev4[0], st, ev_flags), if (ev4) cudaEventRecordWithFlags(ev4[1], st, ev_flags),
                      if (ev4) cudaEventRecordWithFlags(ev4[2], st, ev_flags), if (ev4) cudaEventRecordWithFlags(ev4[3], st, ev_flags));
    CU(cudaGetLastError());
    LAUNCHED(15);
    return CDS_OK;
}

static int launch_wide(PglzJob *jobs, DevState *st_counts, uint32_t n_jobs_host, WideArea wa, int num_sms, cudaStream_t st,
                       bool with_pipeline = true, cudaEvent_t *ev4 = nullptr, unsigned ev_flags = cudaEventRecordDefault,
                       const uint32_t *d_first = nullptr)
{
    k_wide_classify<<<1, 1024, 0, st>>>(jobs, st_counts, n_jobs_host, wa, with_pipeline ? 0 : 1, d_first, d_first ? d_first + 1 : nullptr);
    CU(cudaGetLastError());
    LAUNCHED(1);
    return with_pipeline ? wide_pipeline(wa, num_sms, st, ev4, ev_flags) : CDS_OK;
}

extern "C" int cds_set_wide_pglz_threshold(uint32_t decoded_bytes, int skip_cost_model)
{
    g_wide_threshold = decoded_bytes;
    g_wide_force = skip_cost_model ? 1 : 0;
    return CDS_OK;
}

// ---- handle ------------------------------------------------------------------------------------

template <typename T>
static cudaError_t dalloc(T **p, uint64_t n)
{
    return cudaMalloc(reinterpret_cast<void **>(p), sizeof(T) * (n ? n : 1));
}

extern "C" int cds_segment_destroy(cds_segment *s)
{
    if (!s) return CDS_OK;
    DeviceScope device_scope__(s->device);
    void *bufs[] = {s->stage_heap, s->stage_toast, s->heap_base, s->toast_base, s->rows, s->zero_region, s->scan_vals, s->dir_id, s->dir_loc,
                    s->dir_ref, s->chunk_flag, s->vfirst, s->vcount, s->vseq0, s->vbytes, s->vscr, s->inline_flag,
                    s->row_first, s->pos_of_row, s->row_at_pos, s->ind_len, s->dep_len, s->pos_ind, s->pos_dep,
                    s->row_ind_off, s->row_dep_off, s->ops, s->jobs, s->scratch, s->wide_mem};
    for (void *b : bufs) if (b) cudaFree(b);
    if (s->wide_store) cudaFree(s->wide_store);
    if (s->stream_mem) cudaFree(s->stream_mem);
    if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
    for (cudaStream_t x : s->dec_stream) if (x) cudaStreamDestroy(x);
    for (cudaEvent_t e : s->ev_copy) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : s->ev_batch) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : s->ev_dec) if (e) cudaEventDestroy(e);
    if (s->h_state) cudaFreeHost(s->h_state);
    if (s->d_lut) cudaFree(s->d_lut);
    if (s->ev) {
        for (uint32_t i = 0; i < 2 * kTimedStages * s->ev_cap; i++) cudaEventDestroy(s->ev[i]);
        free(s->ev);
        free(s->ev_mask);
    }
    delete s;
    return CDS_OK;
}

extern "C" int cds_segment_create(cds_segment **out, int device, uint64_t max_heap_pages, uint64_t max_toast_pages,
                                  uint64_t max_chunks, uint64_t scratch_bytes, int stage_pages)
{
    if (!out || max_heap_pages == 0) return CDS_ERR_ARG;
    if (max_heap_pages > 0x7FFFFFFFull / 430 || max_toast_pages > 0x7FFFFFFFull / 8) return CDS_ERR_ARG;
    int ndev = cds_device_count();
    if (ndev == 0) return CDS_ERR_NO_DEVICE;
    if (device < 0 || device >= ndev) return CDS_ERR_ARG;
    ON_DEVICE(device);
    cds_segment *s = new (std::nothrow) cds_segment();
    if (!s) return CDS_ERR_ARG;
    s->device = device;
    cudaDeviceGetAttribute(&s->num_sms, cudaDevAttrMultiProcessorCount, device);
    s->cap_heap_pages = max_heap_pages;
    s->cap_toast_pages = max_toast_pages;
    s->cap_rows = max_heap_pages * 430;   // (32768-24) / (4-byte ItemId + 24+48-byte minimal tuple)
    s->cap_chunks = max_chunks ? max_chunks : max_toast_pages * 6 + 4096;
    s->cap_scratch = scratch_bytes == ~0ull ? max_toast_pages * (uint64_t)kPage : scratch_bytes;
    uint64_t hsize = 1024;
    while (hsize < s->cap_rows * 4) hsize <<= 1;   // 2 values per row, load factor <= 0.5
    s->hmask = (uint32_t)(hsize - 1);
    const uint64_t R = s->cap_rows, V = 2 * R, C = s->cap_chunks;
    cudaError_t e = cudaSuccess;
#define A(ptr, n) if (e == cudaSuccess) e = dalloc(&s->ptr, (n))
    A(heap_base, max_heap_pages + 1); A(toast_base, max_toast_pages + 1); A(rows, R * kRowWords);
    A(dir_id, C); A(dir_loc, C); A(dir_ref, C); A(chunk_flag, C + 1);
    A(vfirst, V); A(vcount, V); A(vseq0, V); A(vbytes, V); A(vscr, V + 1);
    A(inline_flag, R + 1); A(row_first, R); A(pos_of_row, R); A(row_at_pos, R);
    A(ind_len, R); A(dep_len, R); A(pos_ind, R + 1); A(pos_dep, R + 1); A(row_ind_off, R); A(row_dep_off, R);
    A(ops, C); A(jobs, V); A(scratch, s->cap_scratch + 16);
    {
        // scans: [0] heap pages, [1] TOAST pages, [2] rows, [3] chunks, [4],[5] positions, [6] values
        const uint64_t lens[7] = {max_heap_pages, max_toast_pages, R, C, R, R, V};
        uint64_t tiles[7], total_tiles = 0;
        for (int i = 0; i < 7; i++) { tiles[i] = lens[i] / kScanMinTile + 2; total_tiles += tiles[i]; }
        // streamed decode (only with stage buffers): 5 more scans for the assumed plan, one page-index scan per range,
        // StreamState, the batch bookkeeping: all of it starts from zero too
        const uint64_t range_tiles = max_toast_pages / kScanMinTile + 2;
        uint64_t stream_tiles = 0;
        if (stage_pages) stream_tiles = (R / kScanMinTile + 2) * 4 + (V / kScanMinTile + 2) + range_tiles * kMaxStreamRanges;
        const uint64_t stream_words = stage_pages ? 16 + 2 * (kMaxStreamRanges + 2) + 8 + kMaxStreamRanges + stream_tiles : 0;
        const uint64_t state_bytes = 256, ws_bytes = ((8 + total_tiles + stream_words) * 4 + 255) & ~255ull;
        s->zero_bytes = state_bytes + ws_bytes + hsize * 8;
        if (e == cudaSuccess) e = dalloc(&s->zero_region, s->zero_bytes);
        if (e == cudaSuccess) e = dalloc(&s->scan_vals, 2 * (total_tiles + stream_tiles));
        if (e == cudaSuccess) {
            s->state = reinterpret_cast<DevState *>(s->zero_region);
            uint32_t *w = reinterpret_cast<uint32_t *>(s->zero_region + state_bytes);
            s->htab = reinterpret_cast<unsigned long long *>(s->zero_region + state_bytes + ws_bytes);
            uint64_t t0 = 0;
            for (int i = 0; i < 7; i++) {
                s->ws[i].ticket = w + i;
                s->ws[i].status = w + 8 + t0;
                s->ws[i].agg = s->scan_vals + 2 * t0;
                s->ws[i].incl = s->scan_vals + 2 * t0 + tiles[i];
                t0 += tiles[i];
            }
            if (stage_pages) {
                uint32_t *sw = w + 8 + total_tiles;                         // behind the classic scans' status words
                s->sargs.ss = reinterpret_cast<StreamState *>(sw); sw += 16;
                s->sargs.range_first = sw; sw += kMaxStreamRanges + 2;
                s->sargs.range_ticket = sw; sw += kMaxStreamRanges + 2;
                uint32_t *tickets = sw; sw += 8 + kMaxStreamRanges;
                auto carve = [&](ScanWs &x, uint64_t nt, uint32_t *ticket) {
                    x.ticket = ticket; x.status = sw; x.agg = s->scan_vals + 2 * t0; x.incl = s->scan_vals + 2 * t0 + nt;
                    sw += nt; t0 += nt;
                };
                for (int i = 0; i < 4; i++) carve(s->sargs.ws[i], R / kScanMinTile + 2, tickets + i);
                carve(s->sargs.ws[4], V / kScanMinTile + 2, tickets + 4);
                for (uint32_t k = 0; k < kMaxStreamRanges; k++) carve(s->ws_range[k], range_tiles, tickets + 8 + k);
            }
        }
    }
    if (stage_pages) {
        A(stage_heap, max_heap_pages * (uint64_t)kPage); A(stage_toast, max_toast_pages * (uint64_t)kPage);
        // the assumed plan of the streamed decode: [ext_flag R+1 | o_pos_of_row R | vjob V] u32, [o_ind_len R | o_dep_len R] u64
        const uint64_t u32s = (R + 1) + R + V, bytes = ((u32s * 4 + 15) & ~15ull) + 16 * R;
        A(stream_mem, bytes);
        if (e == cudaSuccess) {
            uint32_t *q = reinterpret_cast<uint32_t *>(s->stream_mem);
            s->sargs.ext_flag = q; q += R + 1;
            s->sargs.o_pos_of_row = q; q += R;
            s->sargs.vjob = q;
            unsigned long long *q8 = reinterpret_cast<unsigned long long *>(s->stream_mem + ((u32s * 4 + 15) & ~15ull));
            s->sargs.o_ind_len = q8;
            s->sargs.o_dep_len = q8 + R;
            e = cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking);
            for (uint32_t k = 0; e == cudaSuccess && k < kMaxStreamRanges + 2; k++) e = cudaEventCreateWithFlags(&s->ev_copy[k], cudaEventDisableTiming);
            for (uint32_t k = 0; e == cudaSuccess && k < kMaxStreamRanges + 1; k++) e = cudaEventCreateWithFlags(&s->ev_batch[k], cudaEventDisableTiming);
            for (uint32_t k = 0; e == cudaSuccess && k < kStreamDecoders; k++) {
                e = cudaStreamCreateWithFlags(&s->dec_stream[k], cudaStreamNonBlocking);
                if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->ev_dec[k], cudaEventDisableTiming);
            }
        }
    }
    if (s->cap_scratch) {
        // wide pglz: 8 bytes per 512-byte source segment of the staged (compressed) values
        const uint64_t cj = 1024, cs = s->cap_scratch / kWideSeg + cj * kWideSuper, cb = s->cap_scratch / kWideSeg + cj;
        if (cs < 0xFFFFFFFFull && cb < 0xFFFFFFFFull) {
            A(wide_mem, wide_area_bytes((uint32_t)cj, (uint32_t)cs, (uint32_t)cb));
            if (e == cudaSuccess) s->wide = wide_area_carve(s->wide_mem, (uint32_t)cj, (uint32_t)cs, (uint32_t)cb);
        }
    }
#undef A
    if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void **>(&s->h_state), sizeof(DevState), cudaHostAllocPortable);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_toast_gather, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGatherSmemBytes);
    if (e == cudaSuccess) e = pglz_prepare(device);
    if (e == cudaSuccess) e = wide_prepare();
    if (e == cudaSuccess) {
        int coop = 0, per_sm = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device);
        if (coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_plan_fused, 256, 0) == cudaSuccess && per_sm > 0)
            s->coop_blocks = s->num_sms * (per_sm < 2 ? per_sm : 2);
        cudaGetLastError();
    }
    if (e != cudaSuccess) {
        int rc = cuda_fail(e, "cds_segment_create");
        cds_segment_destroy(s);
        return rc;
    }
    *out = s;
    return CDS_OK;
}

extern "C" int cds_set_output_format(cds_segment *s, int independent_var_format)
{
    if (!s || independent_var_format < 0 || independent_var_format > 2) return CDS_ERR_ARG;
    if (independent_var_format == 2 && !s->d_lut) return CDS_ERR_ARG;       // set the table first
    s->ind_format = independent_var_format;
    return CDS_OK;
}

extern "C" int cds_set_wide_pglz(cds_segment *s, int mode)
{
    if (!s || mode < 0 || mode > 3) return CDS_ERR_ARG;
    s->wide_mode = mode;
    return CDS_OK;
}

extern "C" int cds_set_normalize_lut(cds_segment *s, const float *h_lut, uint32_t channels)
{
    if (!s || !h_lut || channels == 0 || channels > kMaxLutChannels) return CDS_ERR_ARG;
    ON_DEVICE(s->device);
    if (!s->d_lut) CU(cudaMalloc(reinterpret_cast<void **>(&s->d_lut), sizeof(float) * 256 * kMaxLutChannels));
    CU(cudaMemcpy(s->d_lut, h_lut, sizeof(float) * 256 * channels, cudaMemcpyHostToDevice));
    s->lut_channels = channels;
    return CDS_OK;
}

// ---- steps -------------------------------------------------------------------------------------

static inline uint32_t small_grid(const cds_segment *s) { return (uint32_t)s->num_sms * 4u; }
static inline uint32_t scan_grid(const cds_segment *s, uint64_t n, uint32_t items_per_thread)
{
    uint64_t tiles = n / ((uint64_t)kScanThreads * items_per_thread) + 1, cap = (uint64_t)s->num_sms * 4u;
    return (uint32_t)(tiles < cap ? tiles : cap);
}

// A decode starts from zeros: state, scan tickets/status and the hash table, one memset.
static int reset_state(cds_segment *s, cudaStream_t st)
{
    // A streamed cds_decode_host whose verdict was never collected (no cds_finish since): collect it now, so that a segment
    // whose assumed plan did not hold is decoded again before its stage buffers and its state are reused.
    if (s->stream_pending) {
        int rc = cds_finish(s, (void *)st, nullptr);
        if (rc >= 100) return rc;
    }
    CU(cudaMemsetAsync(s->zero_region, 0, s->zero_bytes, st));
    s->have_rows = s->have_chunks = s->have_plan = s->have_ops = false;
    s->wide_launched = false;
    return CDS_OK;
}

extern "C" int cds_heap_scan(cds_segment *s, const void *d_heap_pages, uint64_t n_pages, void *stream)
{
    if (!s || (!d_heap_pages && n_pages)) return CDS_ERR_ARG;
    if (n_pages > s->cap_heap_pages) return CDS_ERR_CAPACITY;
    if (reinterpret_cast<uintptr_t>(d_heap_pages) & 15u) return CDS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(s->device);
    int rc = reset_state(s, st);
    if (rc) return rc;
    s->d_heap = static_cast<const uint8_t *>(d_heap_pages);
    s->n_heap_pages = n_pages;
    k_page_index<<<scan_grid(s, n_pages, 4), kScanThreads, 0, st>>>(s->d_heap, (uint32_t)n_pages, s->heap_base, s->state, 0,
                                                                  (uint32_t)s->cap_rows, s->ws[0]);
    CU(cudaGetLastError());
    LAUNCHED(1);
    if (n_pages) {
        uint64_t want = n_pages * kHeapSub, cap = (uint64_t)s->num_sms * 16u;
        uint32_t grid = (uint32_t)(want < cap ? want : cap);
        k_heap_rows<<<grid, 64, 0, st>>>(s->d_heap, (uint32_t)n_pages, s->heap_base, s->rows, s->state, s->htab, s->hmask,
                                          s->vfirst, s->vcount, s->vbytes, s->vseq0);
        CU(cudaGetLastError());
        LAUNCHED(1);
    }
    s->have_rows = true;
    return CDS_OK;
}

extern "C" int cds_set_rows(cds_segment *s, const uint32_t *h_rows, uint64_t n_rows, const void *d_inline_blob, void *stream)
{
    if (!s || (!h_rows && n_rows)) return CDS_ERR_ARG;
    if (n_rows > s->cap_rows) return CDS_ERR_CAPACITY;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(s->device);
    int rc = reset_state(s, st);
    if (rc) return rc;
    k_set_n_rows<<<1, 1, 0, st>>>(s->state, (uint32_t)n_rows);
    if (n_rows) CU(cudaMemcpyAsync(s->rows, h_rows, n_rows * kRowWords * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));   // h_rows is the caller's (pageable) memory
    s->d_heap = static_cast<const uint8_t *>(d_inline_blob);
    s->n_heap_pages = 0;
    k_index_values<<<small_grid(s), 256, 0, st>>>(s->rows, s->state, s->htab, s->hmask, s->vfirst, s->vcount, s->vbytes, s->vseq0);
    CU(cudaGetLastError());
    LAUNCHED(2);
    s->have_rows = true;
    return CDS_OK;
}

extern "C" int cds_toast_scan(cds_segment *s, const void *d_toast_pages, uint64_t n_pages, void *stream)
{
    if (!s || (!d_toast_pages && n_pages) || !s->have_rows) return CDS_ERR_ARG;
    if (s->have_chunks) return CDS_ERR_ARG;   // one TOAST scan per heap scan: the scan workspaces are single-use
    if (n_pages > s->cap_toast_pages) return CDS_ERR_CAPACITY;
    if (reinterpret_cast<uintptr_t>(d_toast_pages) & 15u) return CDS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(s->device);
    s->d_toast = static_cast<const uint8_t *>(d_toast_pages);
    s->n_toast_pages = n_pages;
    k_page_index<<<scan_grid(s, n_pages, 4), kScanThreads, 0, st>>>(s->d_toast, (uint32_t)n_pages, s->toast_base, s->state, 1,
                                                                  (uint32_t)s->cap_chunks, s->ws[1]);
    CU(cudaGetLastError());
    LAUNCHED(1);
    if (n_pages) {
        k_toast_walk<<<(uint32_t)((n_pages + 127) / 128), 128, 0, st>>>(
            s->d_toast, (uint32_t)n_pages, s->toast_base, s->htab, s->hmask, s->dir_id, s->dir_loc, s->dir_ref,
            s->chunk_flag, s->vfirst, s->vcount, s->vbytes, s->vseq0, s->state, 0u);
        CU(cudaGetLastError());
        LAUNCHED(1);
    }
    s->have_chunks = true;
    return CDS_OK;
}

static PlanArgs plan_args(cds_segment *s, uint32_t isb, uint32_t dsb, void *out_ind, uint64_t ind_cap, void *out_dep,
                          uint64_t dep_cap)
{
    PlanArgs A;
    A.rows = s->rows; A.heap = s->d_heap; A.toast = s->d_toast;
    A.dir_id = s->dir_id; A.dir_loc = s->dir_loc; A.dir_ref = s->dir_ref;
    A.vfirst = s->vfirst; A.vcount = s->vcount; A.vseq0 = s->vseq0; A.vbytes = s->vbytes; A.vscr = s->vscr;
    A.inline_flag = s->inline_flag; A.chunk_flag = s->chunk_flag; A.row_first = s->row_first;
    A.pos_of_row = s->pos_of_row; A.row_at_pos = s->row_at_pos;
    A.ind_len = s->ind_len; A.dep_len = s->dep_len; A.pos_ind = s->pos_ind; A.pos_dep = s->pos_dep;
    A.row_ind_off = s->row_ind_off; A.row_dep_off = s->row_dep_off;
    A.ops = s->ops; A.jobs = s->jobs; A.st = s->state;
    for (int i = 0; i < 5; i++) A.ws[i] = s->ws[2 + i];
    A.isb = isb; A.dsb = dsb;
    A.ind_bf16 = s->ind_format == 1 ? 1u : 0u;
    A.ind_u8norm = s->ind_format == 2 ? 1u : 0u;
    A.lut_channels = s->lut_channels ? s->lut_channels : 1u;
    A.wide_threshold = s->wide_mem ? g_wide_threshold : 0u;
    A.out_ind = static_cast<uint8_t *>(out_ind); A.out_dep = static_cast<uint8_t *>(out_dep); A.scratch = s->scratch;
    A.cap_ind = ind_cap; A.cap_dep = dep_cap; A.cap_scratch = s->cap_scratch;
    return A;
}

// Plan (and, when the output buffers are already known, the copy descriptors) in one cooperative launch;
// separate kernels when cooperative launch is not available.
static int launch_plan(cds_segment *s, PlanArgs &A, bool with_resolve, cudaStream_t st)
{
    if (s->coop_blocks > 0) {
        int do_resolve = with_resolve ? 1 : 0;
        void *args[] = {&A, &do_resolve};
        cudaError_t e = cudaLaunchCooperativeKernel((void *)k_plan_fused, dim3((unsigned)s->coop_blocks), dim3(256), args, 0, st);
        if (e == cudaSuccess) { LAUNCHED(1); return CDS_OK; }
        cudaGetLastError();
        s->coop_blocks = 0;            // fall through to the plain launches from now on
    }
    const uint32_t g = small_grid(s);
    k_plan_rows<<<g, 256, 0, st>>>(A);
    k_plan_scan1<<<dim3(scan_grid(s, s->cap_chunks, 8), 2), kScanThreads, 0, st>>>(A);
    k_assign_pos<<<g, 256, 0, st>>>(A);
    k_plan_scan2<<<dim3(scan_grid(s, 2 * s->cap_rows, 8), 3), kScanThreads, 0, st>>>(A);
    if (with_resolve) k_resolve<<<g, 256, 0, st>>>(A);
    CU(cudaGetLastError());
    LAUNCHED(with_resolve ? 5 : 4);
    return CDS_OK;
}

extern "C" int cds_plan(cds_segment *s, uint32_t ind_sample_bytes, uint32_t dep_sample_bytes, void *stream)
{
    if (!s || !s->have_rows || !s->have_chunks || s->have_plan) return CDS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(s->device);
    PlanArgs A = plan_args(s, ind_sample_bytes, dep_sample_bytes, nullptr, 0, nullptr, 0);
    int rc = launch_plan(s, A, false, st);
    if (rc) return rc;
    s->have_plan = true;
    return CDS_OK;
}

static int launch_gather_and_pglz(cds_segment *s, cudaStream_t st)
{
    const bool timed = s->ev && s->ev_used < s->ev_cap;
    // (inside a stream capture the timing events become event-record nodes of the graph)
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (timed && cudaStreamIsCapturing(st, &cap) != cudaSuccess) { cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
    const unsigned ev_flags = cap == cudaStreamCaptureStatusActive ? cudaEventRecordExternal : cudaEventRecordDefault;
    cudaEvent_t *ev = timed ? s->ev + (size_t)2 * kTimedStages * s->ev_used : nullptr;
    uint8_t mask = 0;
    if (s->n_toast_pages) {
        uint32_t grid = (uint32_t)(s->n_toast_pages < (uint64_t)s->num_sms ? s->n_toast_pages : (uint64_t)s->num_sms);
        if (timed) CU(cudaEventRecordWithFlags(ev[0], st, ev_flags));
        k_toast_gather<<<grid, kGatherThreads, kGatherSmemBytes, st>>>(s->d_toast, (uint32_t)s->n_toast_pages,
                                                                       s->toast_base, s->ops, s->state,
                                                                       s->ind_format == 2 ? s->d_lut : nullptr, s->lut_channels,
                                                                       s->dir_ref, s->vfirst);
        CU(cudaGetLastError());
        LAUNCHED(1);
        if (timed) { CU(cudaEventRecordWithFlags(ev[1], st, ev_flags)); mask |= 1u; }
    }
    const float *lut = s->ind_format == 2 ? s->d_lut : nullptr;
    const bool wide_possible = s->wide_mem && g_wide_threshold && s->wide_mode != 0;
    if (timed) CU(cudaEventRecordWithFlags(ev[2], st, ev_flags));
    if (wide_possible) {
        WideArea wa = s->wide;
        if (s->wide_store_blks * (uint64_t)kWideBlock * sizeof(uint16_t) > g_wide_store_limit) { wa.store = nullptr; wa.store_blks = 0; }
        wa.threshold = g_wide_threshold;
        wa.force = (s->wide_mode == 1 || g_wide_force) ? 1u : 0u;
        wa.st = s->state;
        wa.lut = lut;
        wa.lut_channels = s->lut_channels ? s->lut_channels : 1u;
        s->last_wa = wa;
        // mode 2: the 15 kernels of the wide decoder are enqueued when the handle's last summary saw values go wide;
        // the classification always runs, and what it takes without them is decoded by cds_finish (see there)
        const bool pipeline = s->wide_mode != 2 || s->wide_hint;
        int rc = launch_wide(s->jobs, s->state, 0u, wa, s->num_sms, st, pipeline, (timed && pipeline) ? ev + 4 : nullptr, ev_flags);
        if (rc) return rc;
        if (timed && pipeline) mask |= 4u | 8u;
        s->wide_launched = true;
    }
    // one warp per value; the grid is sized for residency (kPglzCtasPerSm CTAs of kPglzWarps warps fill one SM's shared memory)
    k_pglz_decode<<<(uint32_t)s->num_sms * kPglzCtasPerSm, 32 * kPglzWarps, kPglzSmemBytes, st>>>(
        s->jobs, s->state, 0u, nullptr, s->state, &s->state->pglz_ticket, lut, s->lut_channels);
    CU(cudaGetLastError());
    LAUNCHED(1);
    if (timed) {
        CU(cudaEventRecordWithFlags(ev[3], st, ev_flags));
        mask |= 2u;
        s->ev_mask[s->ev_used] = mask;
        s->ev_used++;
    }
    return CDS_OK;
}

extern "C" int cds_materialize(cds_segment *s, void *d_out_ind, uint64_t ind_cap, void *d_out_dep, uint64_t dep_cap,
                               void *stream)
{
    if (!s || !s->have_plan || s->have_ops) return CDS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(s->device);
    PlanArgs A = plan_args(s, 0, 0, d_out_ind, ind_cap, d_out_dep, dep_cap);
    k_resolve<<<small_grid(s), 256, 0, st>>>(A);
    CU(cudaGetLastError());
    LAUNCHED(1);
    s->have_ops = true;
    return launch_gather_and_pglz(s, st);
}

__global__ void k_clear_flag(DevState *st, uint32_t flag) { st->flags &= ~flag; }

static void decode_key(unsigned long long key, cds_summary *sum)
{
    if (key == kNoErr) { sum->status = CDS_OK; sum->err_stage = 0; sum->err_where = 0; return; }   // err_inv == 0
    sum->status = (uint32_t)(key & 0xFFu);
    sum->err_stage = (uint32_t)(key >> 61);
    sum->err_where = (key >> 28) & 0x1FFFFFFFFull;
}

// The store of resolved blocks follows what the handle's decodes need: when one handed the wide decoder more blocks than the
// store holds (none at first), the store is replaced by one an eighth larger than that, for the decodes that follow --
// if it stays within cds_set_wide_store_limit and half of the device memory that is free.  The stream is idle here.
static void wide_store_follow(cds_segment *s, uint32_t blks)
{
    if (!g_wide_store_limit || blks <= s->wide_store_blks) return;
    const uint64_t want = (uint64_t)blks + blks / 8u + 16u, bytes = want * kWideBlock * sizeof(uint16_t);
    const uint64_t held = s->wide_store_blks * (uint64_t)kWideBlock * sizeof(uint16_t);
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { cudaGetLastError(); return; }
    if (want > 0xFFFFFFFFull || bytes > g_wide_store_limit || bytes > ((uint64_t)free_b + held) / 2u) return;
    if (s->wide_store) cudaFree(s->wide_store);
    s->wide_store = nullptr; s->wide_store_blks = 0;
    if (cudaMalloc(reinterpret_cast<void **>(&s->wide_store), bytes) != cudaSuccess) { cudaGetLastError(); s->wide_store = nullptr; }
    else s->wide_store_blks = want;
    s->wide.store = s->last_wa.store = s->wide_store;
    s->wide.store_blks = s->last_wa.store_blks = (uint32_t)s->wide_store_blks;
}

extern "C" int cds_set_wide_store_limit(uint64_t bytes)
{
    g_wide_store_limit = bytes;
    return CDS_OK;
}

extern "C" int cds_finish(cds_segment *s, void *stream, cds_summary *sum)
{
    if (!s) return CDS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(s->device);
    for (int pass = 0; pass < 3; pass++) {
        CU(cudaMemcpyAsync(s->h_state, s->state, sizeof(DevState), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        const uint32_t flags = s->h_state->flags;
        if (s->stream_pending) {
            // A streamed cds_decode_host: its assumed plan was compared with the complete directory on the device.
            s->stream_pending = false;
            if (flags & (kFlagStreamRedo | kFlagNeedWide)) {
                // It does not hold (or an error was met, or large values turned up whose decoder was not enqueued): decode
                // again the classic way from the pages, which are all in the stage buffers now.
                g_stream_redone++;
                if (flags & kFlagNeedWide) s->wide_hint = true;
                int rc = cds_decode_device(s, s->stage_heap, s->redo.n_heap_pages, s->stage_toast, s->redo.n_toast_pages, s->redo.isb,
                                           s->redo.dsb, s->redo.out_ind, s->redo.ind_cap, s->redo.out_dep, s->redo.dep_cap, stream, nullptr);
                if (rc) return rc;
                continue;
            }
        }
        if (s->wide_launched && (flags & kFlagNeedWide)) {
            // The classification of this decode took values for the wide decoder, but its kernels were not enqueued with
            // the decode (the handle had not seen such values before): decode them now.  One extra round trip, once.
            k_clear_flag<<<1, 1, 0, st>>>(s->state, kFlagNeedWide);
            LAUNCHED(1);
            int rc = wide_pipeline(s->last_wa, s->num_sms, st);
            if (rc) return rc;
            continue;
        }
        break;
    }
    const DevState &h = *s->h_state;
    if (s->wide_launched) {
        g_wide_decoded += h.n_wide_taken;
        s->wide_hint = h.n_wide_taken != 0;
        s->wide_launched = false;
        if (h.n_wide_taken) wide_store_follow(s, h.n_wide_blks);
    }
    cds_summary local;
    memset(&local, 0, sizeof(local));
    local.n_rows = h.n_rows; local.n_positions = h.n_pos; local.ind_bytes = h.ind_bytes; local.dep_bytes = h.dep_bytes;
    local.n_chunks = h.n_chunks; local.n_inline = h.n_inline; local.scratch_bytes = h.scratch_bytes;
    local.n_wide_values = h.n_wide;
    decode_key(~h.err_inv, &local);
    if (sum) *sum = local;
    return (int)local.status;
}

extern "C" int cds_decode_device(cds_segment *s, const void *d_heap, uint64_t n_heap_pages, const void *d_toast,
                                 uint64_t n_toast_pages, uint32_t isb, uint32_t dsb, void *d_out_ind, uint64_t ind_cap,
                                 void *d_out_dep, uint64_t dep_cap, void *stream, cds_summary *sum)
{
    if (!s || (!d_heap && n_heap_pages) || (!d_toast && n_toast_pages)) return CDS_ERR_ARG;
    if (n_heap_pages > s->cap_heap_pages || n_toast_pages > s->cap_toast_pages) return CDS_ERR_CAPACITY;
    if ((reinterpret_cast<uintptr_t>(d_heap) | reinterpret_cast<uintptr_t>(d_toast)) & 15u) return CDS_ERR_ARG;
    int rc;
    ON_DEVICE(s->device);
    {
        // same steps as cds_heap_scan + cds_toast_scan, with both page indexes in one launch
        cudaStream_t st = (cudaStream_t)stream;
        if ((rc = reset_state(s, st))) return rc;
        s->d_heap = static_cast<const uint8_t *>(d_heap); s->n_heap_pages = n_heap_pages;
        s->d_toast = static_cast<const uint8_t *>(d_toast); s->n_toast_pages = n_toast_pages;
        uint32_t gx = scan_grid(s, n_heap_pages > n_toast_pages ? n_heap_pages : n_toast_pages, 4);
        if (n_heap_pages && n_heap_pages <= kFrontMaxHeapPages) {
            // page indexes and the heap row parse in one launch (three independent planes)
            const uint64_t want = (n_heap_pages * kHeapSub + 3) / 4, cap = (uint64_t)s->num_sms * 4u;
            const uint32_t gr = (uint32_t)(want < cap ? want : cap);
            if (gr > gx) gx = gr;
            k_front<<<dim3(gx, 3), kScanThreads, 0, st>>>(s->d_heap, (uint32_t)n_heap_pages, s->heap_base, (uint32_t)s->cap_rows, s->ws[0],
                                                         s->d_toast, (uint32_t)n_toast_pages, s->toast_base, (uint32_t)s->cap_chunks,
                                                         s->ws[1], s->state, s->rows, s->htab, s->hmask, s->vfirst, s->vcount,
                                                         s->vbytes, s->vseq0);
            CU(cudaGetLastError());
            LAUNCHED(1);
        } else {
            k_page_index2<<<dim3(gx, 2), kScanThreads, 0, st>>>(s->d_heap, (uint32_t)n_heap_pages, s->heap_base, (uint32_t)s->cap_rows,
                                                               s->ws[0], s->d_toast, (uint32_t)n_toast_pages, s->toast_base,
                                                               (uint32_t)s->cap_chunks, s->ws[1], s->state);
            CU(cudaGetLastError());
            LAUNCHED(1);
            if (n_heap_pages) {
                uint64_t want = n_heap_pages * kHeapSub, cap = (uint64_t)s->num_sms * 16u;
                k_heap_rows<<<(uint32_t)(want < cap ? want : cap), 64, 0, st>>>(s->d_heap, (uint32_t)n_heap_pages, s->heap_base, s->rows,
                                                                              s->state, s->htab, s->hmask, s->vfirst, s->vcount,
                                                                              s->vbytes, s->vseq0);
                CU(cudaGetLastError());
                LAUNCHED(1);
            }
        }
        if (n_toast_pages) {
            k_toast_walk<<<(uint32_t)((n_toast_pages + 127) / 128), 128, 0, st>>>(
                s->d_toast, (uint32_t)n_toast_pages, s->toast_base, s->htab, s->hmask, s->dir_id, s->dir_loc, s->dir_ref,
                s->chunk_flag, s->vfirst, s->vcount, s->vbytes, s->vseq0, s->state, 0u);
            CU(cudaGetLastError());
            LAUNCHED(1);
        }
        s->have_rows = s->have_chunks = true;
    }
    {
        cudaStream_t st = (cudaStream_t)stream;
        PlanArgs A = plan_args(s, isb, dsb, d_out_ind, ind_cap, d_out_dep, dep_cap);
        if ((rc = launch_plan(s, A, true, st))) return rc;
        s->have_plan = s->have_ops = true;
        if ((rc = launch_gather_and_pglz(s, st))) return rc;
    }
    return sum ? cds_finish(s, stream, sum) : CDS_OK;
}

// The pglz stage over batch j of the streamed decode (slots [range_first[j], range_first[j + 1]) of the plain job list).
// The classification and the wide decoder share one work area, so they stay on the main stream; the warp decoder of the
// batch runs on a side stream, beside the ranges that follow (a 150 KB value keeps its warp busy for milliseconds: batches
// must overlap each other, or the decode falls behind the copy).
static int launch_pglz_batch(cds_segment *s, uint32_t j, bool may_be_wide, cudaStream_t st)
{
    const float *lut = s->ind_format == 2 ? s->d_lut : nullptr;
    const uint32_t *d_first = s->sargs.range_first + j;
    if (may_be_wide && s->wide_mem && g_wide_threshold && s->wide_mode != 0) {
        WideArea wa = s->wide;
        if (s->wide_store_blks * (uint64_t)kWideBlock * sizeof(uint16_t) > g_wide_store_limit) { wa.store = nullptr; wa.store_blks = 0; }
        wa.threshold = g_wide_threshold;
        wa.force = (s->wide_mode == 1 || g_wide_force) ? 1u : 0u;
        wa.st = s->state;
        wa.lut = lut;
        wa.lut_channels = s->lut_channels ? s->lut_channels : 1u;
        s->last_wa = wa;
        const bool pipeline = s->wide_mode != 2 || s->wide_hint;
        int rc = launch_wide(s->jobs, s->state, 0u, wa, s->num_sms, st, pipeline, nullptr, cudaEventRecordDefault, d_first);
        if (rc) return rc;
        s->wide_launched = true;
    }
    cudaStream_t side = s->dec_stream[j % kStreamDecoders];
    CU(cudaEventRecord(s->ev_batch[j], st));
    CU(cudaStreamWaitEvent(side, s->ev_batch[j], 0));
    k_pglz_decode<<<(uint32_t)s->num_sms * kPglzCtasPerSm, 32 * kPglzWarps, kPglzSmemBytes, side>>>(
        s->jobs, s->state, 0u, nullptr, s->state, s->sargs.range_ticket + j, lut, s->lut_channels, d_first, d_first + 1);
    CU(cudaGetLastError());
    LAUNCHED(1);
    return CDS_OK;
}

// cds_decode_host with the decode overlapped with the copy (stream_decode.cuh).  The heap goes first; the TOAST relation
// follows in ranges of `stream_range_pages` pages on a copy stream, and each range is processed on `st` as soon as its
// event has fired.  The check of the assumed plan is enqueued last; cds_finish reads its verdict.
static int decode_host_streamed(cds_segment *s, const void *h_heap, uint64_t n_heap_pages, const void *h_toast, uint64_t n_toast_pages,
                                uint32_t isb, uint32_t dsb, void *d_out_ind, uint64_t ind_cap, void *d_out_dep, uint64_t dep_cap,
                                cudaStream_t st)
{
    const uint64_t rp = s->stream_range_pages & 0x7FFFFFFFu;
    const uint32_t n_ranges = (uint32_t)((n_toast_pages + rp - 1) / rp);
    int rc;
    if ((rc = reset_state(s, st))) return rc;
    // the stage buffers may still be read by work enqueued earlier on `st`
    CU(cudaEventRecord(s->ev_copy[kMaxStreamRanges + 1], st));
    CU(cudaStreamWaitEvent(s->copy_stream, s->ev_copy[kMaxStreamRanges + 1], 0));
    CU(cudaMemcpyAsync(s->stage_heap, h_heap, n_heap_pages * (uint64_t)kPage, cudaMemcpyHostToDevice, s->copy_stream));
    CU(cudaEventRecord(s->ev_copy[kMaxStreamRanges], s->copy_stream));
    for (uint32_t k = 0; k < n_ranges; k++) {
        const uint64_t p0 = k * rp, np = (p0 + rp <= n_toast_pages ? rp : n_toast_pages - p0);
        CU(cudaMemcpyAsync(s->stage_toast + p0 * kPage, static_cast<const uint8_t *>(h_toast) + p0 * kPage, np * (uint64_t)kPage,
                           cudaMemcpyHostToDevice, s->copy_stream));
        CU(cudaEventRecord(s->ev_copy[k], s->copy_stream));
    }
    s->d_heap = s->stage_heap; s->n_heap_pages = n_heap_pages;
    s->d_toast = s->stage_toast; s->n_toast_pages = n_toast_pages;
    // ---- heap: rows, value index, and the plan the heap alone implies
    CU(cudaStreamWaitEvent(st, s->ev_copy[kMaxStreamRanges], 0));
    k_page_index<<<scan_grid(s, n_heap_pages, 4), kScanThreads, 0, st>>>(s->d_heap, (uint32_t)n_heap_pages, s->heap_base, s->state, 0,
                                                                       (uint32_t)s->cap_rows, s->ws[0]);
    {
        uint64_t want = n_heap_pages * kHeapSub, cap = (uint64_t)s->num_sms * 16u;
        k_heap_rows<<<(uint32_t)(want < cap ? want : cap), 64, 0, st>>>(s->d_heap, (uint32_t)n_heap_pages, s->heap_base, s->rows, s->state,
                                                                      s->htab, s->hmask, s->vfirst, s->vcount, s->vbytes, s->vseq0);
    }
    PlanArgs A = plan_args(s, isb, dsb, d_out_ind, ind_cap, d_out_dep, dep_cap);
    const StreamArgs S = s->sargs;
    const uint32_t g = small_grid(s);
    k_stream_rows<<<g, 256, 0, st>>>(A, S);
    k_stream_scan1<<<dim3(scan_grid(s, s->cap_rows, 8), 2), kScanThreads, 0, st>>>(A, S);
    k_stream_assign<<<g, 256, 0, st>>>(A, S);
    k_stream_scan2<<<dim3(scan_grid(s, 2 * s->cap_rows, 8), 3), kScanThreads, 0, st>>>(A, S);
    k_stream_mark<<<1, 1, 0, st>>>(s->state, S.range_first);
    k_stream_inline_jobs<<<g, 256, 0, st>>>(A, S);
    k_stream_mark<<<1, 1, 0, st>>>(s->state, S.range_first + 1);
    CU(cudaGetLastError());
    LAUNCHED(9);
    if ((rc = launch_pglz_batch(s, 0, false, st))) return rc;
    // ---- TOAST, range by range
    for (uint32_t k = 0; k < n_ranges; k++) {
        const uint64_t p0 = k * rp, np = (p0 + rp <= n_toast_pages ? rp : n_toast_pages - p0);
        const uint8_t *pages = s->d_toast + p0 * kPage;
        uint32_t *base = s->toast_base + p0;
        CU(cudaStreamWaitEvent(st, s->ev_copy[k], 0));
        k_page_index<<<scan_grid(s, np, 4), kScanThreads, 0, st>>>(pages, (uint32_t)np, base, s->state, 1, (uint32_t)s->cap_chunks, s->ws_range[k]);
        k_stream_fixup<<<1, 256, 0, st>>>(base, (uint32_t)np, s->state, S.ss, (uint32_t)s->cap_chunks);
        k_toast_walk<<<(uint32_t)((np + 127) / 128), 128, 0, st>>>(pages, (uint32_t)np, base, s->htab, s->hmask, s->dir_id, s->dir_loc, s->dir_ref,
                                                                  s->chunk_flag, s->vfirst, s->vcount, s->vbytes, s->vseq0, s->state, (uint32_t)p0);
        k_stream_resolve<<<g, 256, 0, st>>>(A, S, base, base + np);
        k_stream_mark<<<1, 1, 0, st>>>(s->state, S.range_first + k + 2);
        const uint32_t grid = (uint32_t)(np < (uint64_t)s->num_sms ? np : (uint64_t)s->num_sms);
        k_toast_gather<<<grid, kGatherThreads, kGatherSmemBytes, st>>>(pages, (uint32_t)np, base, s->ops, s->state,
                                                                       s->ind_format == 2 ? s->d_lut : nullptr, s->lut_channels, s->dir_ref, s->vfirst);
        CU(cudaGetLastError());
        LAUNCHED(6);
        if ((rc = launch_pglz_batch(s, k + 1, true, st))) return rc;
    }
    // the decoders of all batches join `st` here: what follows rebuilds the job list they read, and whatever the caller
    // enqueues on `st` next must see the complete decode
    for (uint32_t i = 0; i < kStreamDecoders; i++) {
        CU(cudaEventRecord(s->ev_dec[i], s->dec_stream[i]));
        CU(cudaStreamWaitEvent(st, s->ev_dec[i], 0));
    }
    // ---- the classic plan and checks over the complete directory, compared with what was assumed
    k_stream_prefinal<<<1, 1, 0, st>>>(s->state);
    s->have_rows = s->have_chunks = true;
    if ((rc = launch_plan(s, A, true, st))) return rc;
    k_stream_compare<<<g, 256, 0, st>>>(A, S);
    CU(cudaGetLastError());
    LAUNCHED(2);
    s->have_plan = s->have_ops = true;
    s->stream_pending = true;
    s->redo = {n_heap_pages, n_toast_pages, isb, dsb, d_out_ind, d_out_dep, ind_cap, dep_cap};
    g_streamed++;
    return CDS_OK;
}

extern "C" int cds_decode_host(cds_segment *s, const void *h_heap, uint64_t n_heap_pages, const void *h_toast,
                               uint64_t n_toast_pages, uint32_t isb, uint32_t dsb, void *d_out_ind, uint64_t ind_cap,
                               void *d_out_dep, uint64_t dep_cap, void *stream, cds_summary *sum)
{
    if (!s || !s->stage_heap || !s->stage_toast) return CDS_ERR_ARG;
    if ((!h_heap && n_heap_pages) || (!h_toast && n_toast_pages)) return CDS_ERR_ARG;
    if (n_heap_pages > s->cap_heap_pages || n_toast_pages > s->cap_toast_pages) return CDS_ERR_CAPACITY;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(s->device);
    // Streamed when the TOAST relation spans more than three ranges (below that the copy is too short to hide anything: a
    // Criteo segment of 104 MiB copies in 2 ms and decodes in 10) and the calling stream is not being captured (the copy
    // stream and its events are the library's own).
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cap) != cudaSuccess) { cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
    const uint64_t rp = s->stream_range_pages & 0x7FFFFFFFu;
    const uint64_t min_ranges = (s->stream_range_pages & 0x80000000u) ? 1 : 3;   // (tests: stream anything of more than one range)
    if (rp && n_heap_pages && n_toast_pages > min_ranges * rp && (n_toast_pages + rp - 1) / rp <= kMaxStreamRanges && cap == cudaStreamCaptureStatusNone &&
        !(s->ev && s->ev_used < s->ev_cap)) {
        int rc = decode_host_streamed(s, h_heap, n_heap_pages, h_toast, n_toast_pages, isb, dsb, d_out_ind, ind_cap, d_out_dep, dep_cap, st);
        if (rc) return rc;
        return sum ? cds_finish(s, stream, sum) : CDS_OK;
    }
    CU(cudaMemcpyAsync(s->stage_heap, h_heap, n_heap_pages * (uint64_t)kPage, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s->stage_toast, h_toast, n_toast_pages * (uint64_t)kPage, cudaMemcpyHostToDevice, st));
    return cds_decode_device(s, s->stage_heap, n_heap_pages, s->stage_toast, n_toast_pages, isb, dsb, d_out_ind, ind_cap,
                             d_out_dep, dep_cap, stream, sum);
}

extern "C" int cds_set_host_streaming(cds_segment *s, uint32_t range_pages)
{
    if (!s) return CDS_ERR_ARG;
    s->stream_range_pages = range_pages;
    return CDS_OK;
}

extern "C" int cds_stream_counts(unsigned long long *streamed, unsigned long long *decoded_again)
{
    if (streamed) *streamed = g_streamed;
    if (decoded_again) *decoded_again = g_stream_redone;
    return CDS_OK;
}

// ---- measurement hooks ---------------------------------------------------------------------------

extern "C" unsigned long long cds_launch_count(void) { return g_launches; }
extern "C" unsigned long long cds_wide_decoded_count(void) { return g_wide_decoded; }

extern "C" int cds_timing_enable(cds_segment *s, uint32_t max_samples)
{
    if (!s) return CDS_ERR_ARG;
    ON_DEVICE(s->device);
    if (s->ev) {
        for (uint32_t i = 0; i < 2 * kTimedStages * s->ev_cap; i++) cudaEventDestroy(s->ev[i]);
        free(s->ev);
        free(s->ev_mask);
        s->ev = nullptr;
        s->ev_mask = nullptr;
    }
    s->ev_cap = s->ev_used = 0;
    if (max_samples == 0) return CDS_OK;
    s->ev = (cudaEvent_t *)calloc(2 * kTimedStages * (size_t)max_samples, sizeof(cudaEvent_t));
    s->ev_mask = (uint8_t *)calloc(max_samples, 1);
    if (!s->ev || !s->ev_mask) return CDS_ERR_ARG;
    for (uint32_t i = 0; i < 2 * kTimedStages * max_samples; i++) CU(cudaEventCreate(&s->ev[i]));
    s->ev_cap = max_samples;
    return CDS_OK;
}

static int timing_read(cds_segment *s, uint32_t stage, float *h_ms, uint32_t cap, uint32_t *n, bool rearm)
{
    if (!s || !n || stage >= kTimedStages) return CDS_ERR_ARG;
    ON_DEVICE(s->device);
    uint32_t k = 0;
    for (uint32_t i = 0; i < s->ev_used && k < cap; i++) {
        if (!((s->ev_mask[i] >> stage) & 1u)) continue;
        cudaEvent_t *ev = s->ev + (size_t)2 * kTimedStages * i + 2 * stage;
        CU(cudaEventSynchronize(ev[1]));
        CU(cudaEventElapsedTime(&h_ms[k], ev[0], ev[1]));
        k++;
    }
    *n = k;
    if (rearm) s->ev_used = 0;
    return CDS_OK;
}

extern "C" int cds_timing_read(cds_segment *s, float *h_ms, uint32_t cap, uint32_t *n) { return timing_read(s, 0, h_ms, cap, n, true); }

extern "C" int cds_timing_read_stage(cds_segment *s, uint32_t stage, float *h_ms, uint32_t cap, uint32_t *n)
{
    return timing_read(s, stage, h_ms, cap, n, false);
}

// ---- exports for the Python mirror -------------------------------------------------------------

extern "C" int cds_export_rows(cds_segment *s, uint32_t *h_rows, uint64_t cap_rows, uint64_t *n_rows, void *stream)
{
    if (!s || !n_rows || !s->have_rows) return CDS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(s->device);
    CU(cudaMemcpyAsync(s->h_state, s->state, sizeof(DevState), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    *n_rows = s->h_state->n_rows;
    if (*n_rows > cap_rows) return CDS_ERR_CAPACITY;
    if (*n_rows && h_rows) CU(cudaMemcpy(h_rows, s->rows, *n_rows * kRowWords * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    return CDS_OK;
}

extern "C" int cds_export_layout(cds_segment *s, uint64_t *h_order, uint64_t *h_ind_off, uint64_t *h_dep_off,
                                 uint64_t *h_ind_len, uint64_t *h_dep_len, uint64_t cap_rows, void *stream)
{
    if (!s || !s->have_plan) return CDS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(s->device);
    CU(cudaMemcpyAsync(s->h_state, s->state, sizeof(DevState), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const uint64_t n_rows = s->h_state->n_rows, n_pos = s->h_state->n_pos;
    if (n_rows > cap_rows) return CDS_ERR_CAPACITY;
    if (n_pos && h_order) {
        uint32_t *tmp = (uint32_t *)malloc(sizeof(uint32_t) * n_pos);
        if (!tmp) return CDS_ERR_ARG;
        cudaError_t e = cudaMemcpy(tmp, s->row_at_pos, sizeof(uint32_t) * n_pos, cudaMemcpyDeviceToHost);
        for (uint64_t i = 0; e == cudaSuccess && i < n_pos; i++) h_order[i] = tmp[i];
        free(tmp);
        if (e != cudaSuccess) return cuda_fail(e, "export order");
    }
    if (n_rows) {
        if (h_ind_off) CU(cudaMemcpy(h_ind_off, s->row_ind_off, 8 * n_rows, cudaMemcpyDeviceToHost));
        if (h_dep_off) CU(cudaMemcpy(h_dep_off, s->row_dep_off, 8 * n_rows, cudaMemcpyDeviceToHost));
        if (h_ind_len) CU(cudaMemcpy(h_ind_len, s->ind_len, 8 * n_rows, cudaMemcpyDeviceToHost));
        if (h_dep_len) CU(cudaMemcpy(h_dep_len, s->dep_len, 8 * n_rows, cudaMemcpyDeviceToHost));
    }
    return CDS_OK;
}

extern "C" int cds_export_chunks(cds_segment *s, uint32_t *h_id, uint32_t *h_seq, uint64_t *h_off, uint32_t *h_len,
                                 uint64_t cap_chunks, uint64_t *n_chunks, void *stream)
{
    if (!s || !n_chunks || !s->have_chunks) return CDS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(s->device);
    CU(cudaMemcpyAsync(s->h_state, s->state, sizeof(DevState), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const uint64_t n = s->h_state->n_chunks;
    *n_chunks = n;
    if (n > cap_chunks) return CDS_ERR_CAPACITY;
    if (n == 0) return CDS_OK;
    uint2 *a = (uint2 *)malloc(sizeof(uint2) * n), *b = (uint2 *)malloc(sizeof(uint2) * n);
    if (!a || !b) { free(a); free(b); return CDS_ERR_ARG; }
    cudaError_t e = cudaMemcpy(a, s->dir_id, sizeof(uint2) * n, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(b, s->dir_loc, sizeof(uint2) * n, cudaMemcpyDeviceToHost);
    for (uint64_t i = 0; e == cudaSuccess && i < n; i++) {
        if (h_id) h_id[i] = a[i].x;
        if (h_seq) h_seq[i] = a[i].y;
        if (h_off) h_off[i] = (uint64_t)b[i].x * kPage + (b[i].y & 0xFFFFu);
        if (h_len) h_len[i] = b[i].y >> 16;
    }
    free(a); free(b);
    if (e != cudaSuccess) return cuda_fail(e, "export chunks");
    return CDS_OK;
}

// ---- pglz on its own ---------------------------------------------------------------------------

// One thread per datum: read the 8-byte header on the device and fill the job slot.
__global__ void __launch_bounds__(256)
k_build_pglz_jobs(const uint8_t *__restrict__ src, const unsigned long long *__restrict__ src_off, uint8_t *dst,
                  const unsigned long long *__restrict__ dst_off, uint32_t n, PglzJob *__restrict__ jobs,
                  int32_t *__restrict__ status, DevState *st)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        PglzJob j;
        j.src = 0; j.dst = 0; j.slen = kNone; j.rawsize = 0;
        j.err_key = make_err_key(2, i, 0, 0, CDS_ERR_PGLZ_CORRUPT);
        const unsigned long long avail = src_off[i + 1] - src_off[i], room = dst_off[i + 1] - dst_off[i];
        int32_t stat = CDS_ERR_PGLZ_CORRUPT;
        if (avail >= 8) {
            const uint8_t *h = src + src_off[i];
            const uint32_t varsize = ld_u32be(h) & 0x3FFFFFFFu, rawsize = ld_u32(h + 4);
            if (varsize >= 8 && varsize <= avail && rawsize <= room) {
                j.src = (unsigned long long)(h + 8);
                j.dst = (unsigned long long)(dst + dst_off[i]);
                j.slen = varsize - 8;
                j.rawsize = rawsize;
                stat = CDS_OK;
            }
        }
        if (stat != CDS_OK && st) report_key(st, j.err_key);
        status[i] = stat;
        jobs[i] = j;
    }
}

// Work memory of the synchronous helper entry points (cds_pglz_decompress_batch, pglz_decompress, cds_text_etl):
// one grow-only device allocation per (device, purpose), kept for the life of the process (cds_trim_work_memory
// frees them), so that a steady stream of calls never reaches cudaMalloc / cudaFree.  The calls are synchronous,
// so one lock per purpose serialises them.
struct WorkArea {
    uint8_t *ptr = nullptr;
    size_t cap = 0;
};
enum { kWorkBatch = 0, kWorkText = 1, kWorkDatum = 2, kWorkKinds = 3 };
static std::mutex g_work_mu[kWorkKinds];
static WorkArea g_work[64][kWorkKinds];

static cudaError_t work_reserve(int device, int kind, size_t bytes, uint8_t **out)
{
    WorkArea &w = g_work[device][kind];
    if (bytes > w.cap) {
        if (w.ptr) cudaFree(w.ptr);
        w.ptr = nullptr;
        w.cap = 0;
        size_t want = bytes + bytes / 4 + 4096;                               // headroom: sizes that creep up do not reallocate every call
        cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&w.ptr), want);
        if (e != cudaSuccess) { w.ptr = nullptr; return e; }
        w.cap = want;
    }
    *out = w.ptr;
    return cudaSuccess;
}

extern "C" int cds_trim_work_memory(int device)
{
    if (device < 0 || device >= 64) return CDS_ERR_ARG;
    if (cds_device_count() == 0) return CDS_ERR_NO_DEVICE;
    ON_DEVICE(device);
    for (int k = 0; k < kWorkKinds; k++) {
        std::lock_guard<std::mutex> lock(g_work_mu[k]);
        WorkArea &w = g_work[device][k];
        if (w.ptr) cudaFree(w.ptr);
        w.ptr = nullptr;
        w.cap = 0;
    }
    return CDS_OK;
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

extern "C" int cds_pglz_decompress_batch(int device, const void *d_src, const uint64_t *h_src_off, void *d_dst,
                                         const uint64_t *h_dst_off, uint64_t n, int32_t *h_status, void *stream)
{
    if (n == 0) return CDS_OK;
    if (!d_src || !h_src_off || !d_dst || !h_dst_off || n > 0x7FFFFFFFull) return CDS_ERR_ARG;
    if (cds_device_count() == 0) return CDS_ERR_NO_DEVICE;
    if (device < 0 || device >= 64) return CDS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(device);
    // values large enough for the intra-value parallel decoder?  (room bounds rawsize from above)
    bool any_wide = false;
    const uint32_t thr = g_wide_threshold;
    for (uint64_t i = 0; thr && i < n && !any_wide; i++) any_wide = h_dst_off[i + 1] - h_dst_off[i] >= (thr > kWideMinRaw ? thr : kWideMinRaw);
    uint32_t wj = 0, ws_ = 0, wb = 0;
    if (any_wide) {
        const uint64_t src_bytes = h_src_off[n] - h_src_off[0], dst_bytes = h_dst_off[n] - h_dst_off[0];
        const uint64_t cj = n < 65536 ? n : 65536, cs = src_bytes / kWideSeg + (cj + 1) * kWideSuper, cb = dst_bytes / kWideBlock + cj + 1;
        if (cs < 0xFFFFFFFFull && cb < 0xFFFFFFFFull) { wj = (uint32_t)cj; ws_ = (uint32_t)cs; wb = (uint32_t)cb; }
        else any_wide = false;
    }
    // work area of this call: [offset tables | jobs | status words + ticket | wide area]
    const size_t off_bytes = align256(sizeof(uint64_t) * 2 * (n + 1)), job_bytes = align256(sizeof(PglzJob) * n),
                 stat_bytes = align256(sizeof(int32_t) * (n + 1));
    const size_t total = off_bytes + job_bytes + stat_bytes + (any_wide ? align256(wide_area_bytes(wj, ws_, wb)) : 0);
    // ... | store of resolved blocks] when the call has wide values and the store fits (else they are parsed twice)
    size_t store_bytes = any_wide ? (size_t)wb * kWideBlock * sizeof(uint16_t) : 0;
    if (store_bytes > g_wide_store_limit) store_bytes = 0;
    std::lock_guard<std::mutex> lock(g_work_mu[kWorkBatch]);
    uint8_t *base = nullptr;
    cudaError_t e = work_reserve(device, kWorkBatch, total + store_bytes, &base);
    if (e != cudaSuccess && store_bytes) { cudaGetLastError(); store_bytes = 0; e = work_reserve(device, kWorkBatch, total, &base); }
    if (e != cudaSuccess) return cuda_fail(e, "cds_pglz_decompress_batch: work area");
    unsigned long long *d_off = reinterpret_cast<unsigned long long *>(base);
    PglzJob *d_jobs = reinterpret_cast<PglzJob *>(base + off_bytes);
    int32_t *d_stat = reinterpret_cast<int32_t *>(base + off_bytes + job_bytes);
    uint8_t *d_wide = base + off_bytes + job_bytes + stat_bytes;
    e = cudaMemsetAsync(d_stat + n, 0, sizeof(int32_t), st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_off, h_src_off, sizeof(uint64_t) * (n + 1), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_off + n + 1, h_dst_off, sizeof(uint64_t) * (n + 1), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && any_wide) e = wide_prepare();
    if (e == cudaSuccess) e = pglz_prepare(device);
    if (e == cudaSuccess) {
        const int sms = sm_count(device);
        k_build_pglz_jobs<<<(uint32_t)((n + 255) / 256), 256, 0, st>>>(static_cast<const uint8_t *>(d_src), d_off,
                                                                       static_cast<uint8_t *>(d_dst), d_off + n + 1, (uint32_t)n,
                                                                       d_jobs, d_stat, nullptr);
        if (any_wide) {
            WideArea wa = wide_area_carve(d_wide, wj, ws_, wb);
            wa.threshold = thr;
            wa.force = g_wide_force ? 1u : 0u;
            wa.status = d_stat;
            if (store_bytes) { wa.store = reinterpret_cast<uint16_t *>(base + total); wa.store_blks = wb; }
            if (cudaMemsetAsync(wa.state, 0, 256, st) != cudaSuccess || launch_wide(d_jobs, nullptr, (uint32_t)n, wa, sms, st) != CDS_OK) e = cudaErrorUnknown;
        }
        uint32_t grid = (uint32_t)((n + kPglzWarps - 1) / kPglzWarps);
        if (grid > (uint32_t)sms * kPglzCtasPerSm) grid = (uint32_t)sms * kPglzCtasPerSm;
        k_pglz_decode<<<grid, 32 * kPglzWarps, kPglzSmemBytes, st>>>(d_jobs, nullptr, (uint32_t)n, d_stat, nullptr,
                                                                     reinterpret_cast<uint32_t *>(d_stat + n), nullptr, 0u);
        if (e == cudaSuccess) e = cudaGetLastError();
        LAUNCHED(2);
    }
    if (e == cudaSuccess && h_status) e = cudaMemcpyAsync(h_status, d_stat, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, st);
    WideState wc;
    memset(&wc, 0, sizeof(wc));
    if (e == cudaSuccess && any_wide) e = cudaMemcpyAsync(&wc, d_wide, sizeof(wc), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return cuda_fail(e, "cds_pglz_decompress_batch");
    g_wide_decoded += wc.n_taken;
    return CDS_OK;
}

// ---- pglz batch for callers that must not allocate ---------------------------------------------
//
// Some hosts of this library cannot call cudaMalloc on the decode path: a data-loader thread that
// runs while a training framework owns the whole device memory pool, a latency-sensitive ingest
// loop, or a process under a fixed memory budget where an allocation may fail (or synchronise the
// device) at an arbitrary moment.
// cds_pglz_decompress_batch needs two work arrays (job descriptors, status words) and gets them
// from cudaMalloc on every call.  The variant below keeps its work area in a fixed array that is
// part of the loaded module instead, so it never asks the CUDA allocator for anything.
struct StaticScratch {
    DevState state;                          // first-error word; must be zero when a call starts
    PglzJob jobs[CDS_STATIC_MAX_DATUMS];     // rewritten for [0, n) by every call
};
// One copy of this array exists on every GPU that loads the library, from load until exit
// (about 128 KB).  On the host side the NAME below is only a handle for the runtime: it is not a
// device address and must never be handed to a kernel.  static_scratch_for() asks the runtime for
// the real address, per device.
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

static std::mutex g_static_mu;
static void *g_static_addr[64] = {nullptr};

// Device address of g_static_scratch on `device` (which must be the current device).  The first call
// for a device performs the look-up; later calls return the cached address.
static int static_scratch_for(int device, StaticScratch **out)
{
    if (device < 0 || device >= 64) return CDS_ERR_ARG;
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_addr[device]) {
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_static_scratch)");
        if (!ptr) return CDS_ERR_CUDA;
        g_static_addr[device] = ptr;
    }
    *out = static_cast<StaticScratch *>(g_static_addr[device]);
    return CDS_OK;
}

__global__ void k_publish_first_corrupt(const DevState *st, long long *first_corrupt)
{
    const unsigned long long key = ~st->err_inv;
    *first_corrupt = st->err_inv == 0 ? -1ll : (long long)((key >> 28) & 0x1FFFFFFFFull);
}

extern "C" int cds_pglz_decompress_batch_static(int device, const void *d_src, const uint64_t *d_src_off, void *d_dst,
                                                const uint64_t *d_dst_off, uint64_t n, int32_t *d_status,
                                                int64_t *d_first_corrupt, void *stream)
{
    if (!d_src || !d_src_off || !d_dst || !d_dst_off || !d_status) return CDS_ERR_ARG;
    if (n > CDS_STATIC_MAX_DATUMS) {
        snprintf(g_last_error, sizeof(g_last_error),
                 "cds_pglz_decompress_batch_static: %llu datums requested, the built-in work area holds at most %u "
                 "(CDS_STATIC_MAX_DATUMS); split the batch or use cds_pglz_decompress_batch",
                 (unsigned long long)n, (unsigned)CDS_STATIC_MAX_DATUMS);
        return CDS_ERR_CAPACITY;
    }
    if (cds_device_count() == 0) return CDS_ERR_NO_DEVICE;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(device);
    StaticScratch *sc = nullptr;
    int rc = static_scratch_for(device, &sc);
    if (rc != CDS_OK) return rc;
    CU(pglz_prepare(device));                                   // (a function attribute, set once per device: no allocation)
    // The work area still holds whatever the previous call left: the error word is the only part that is
    // accumulated into (atomicMax), so it is the only part that has to be cleared.
    CU(cudaMemsetAsync(&sc->state, 0, sizeof(DevState), st));
    if (n) {
        k_build_pglz_jobs<<<(uint32_t)((n + 255) / 256), 256, 0, st>>>(
            static_cast<const uint8_t *>(d_src), reinterpret_cast<const unsigned long long *>(d_src_off),
            static_cast<uint8_t *>(d_dst), reinterpret_cast<const unsigned long long *>(d_dst_off), (uint32_t)n, sc->jobs,
            d_status, &sc->state);
        CU(cudaGetLastError());
        uint32_t grid = (uint32_t)((n + kPglzWarps - 1) / kPglzWarps);
        const uint32_t max_grid = (uint32_t)sm_count(device) * kPglzCtasPerSm;
        if (grid > max_grid) grid = max_grid;
        k_pglz_decode<<<grid, 32 * kPglzWarps, kPglzSmemBytes, st>>>(sc->jobs, nullptr, (uint32_t)n, d_status, &sc->state,
                                                                     &sc->state.pglz_ticket, nullptr, 0u);
        CU(cudaGetLastError());
        LAUNCHED(2);
    }
    if (d_first_corrupt) {
        k_publish_first_corrupt<<<1, 1, 0, st>>>(&sc->state, reinterpret_cast<long long *>(d_first_corrupt));
        CU(cudaGetLastError());
        LAUNCHED(1);
    }
    return CDS_OK;
}

extern "C" void pglz_decompress(const void *source, char *dest)
{
    // Same contract as pg_lzcompress.c:49-50: the caller sized `dest` from the rawsize field.
    const uint8_t *h = static_cast<const uint8_t *>(source);
    const uint32_t varsize = (((uint32_t)h[0] << 24) | ((uint32_t)h[1] << 16) | ((uint32_t)h[2] << 8) | h[3]) & 0x3FFFFFFFu;
    const uint32_t rawsize = (uint32_t)h[4] | ((uint32_t)h[5] << 8) | ((uint32_t)h[6] << 16) | ((uint32_t)h[7] << 24);
    int dev = 0;
    if (cds_device_count() == 0 || cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        fprintf(stderr, "pglz_decompress: no CUDA device (libcds_da has no CPU path)\n");
        abort();
    }
    // staging buffers of this entry point: grow-only, shared by all callers (the reference calls it from one thread)
    const size_t src_bytes = align256((size_t)(varsize < 8 ? 8 : varsize) + 16), dst_bytes = align256((size_t)rawsize + 16);
    int32_t status = CDS_ERR_CUDA;
    std::lock_guard<std::mutex> lock(g_work_mu[kWorkDatum]);
    uint8_t *base = nullptr;
    cudaError_t e = work_reserve(dev, kWorkDatum, src_bytes + dst_bytes, &base);
    uint8_t *d_src = base, *d_dst = base + src_bytes;
    if (e == cudaSuccess) e = cudaMemcpy(d_src, h, varsize < 8 ? 8 : varsize, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        uint64_t so[2] = {0, varsize < 8 ? 8u : varsize}, dof[2] = {0, rawsize};
        int rc = cds_pglz_decompress_batch(dev, d_src, so, d_dst, dof, 1, &status, nullptr);
        if (rc != CDS_OK) status = rc;
    }
    if (e == cudaSuccess && rawsize) e = cudaMemcpy(dest, d_dst, rawsize, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) {
        fprintf(stderr, "pglz_decompress: %s\n", cudaGetErrorString(e));
        abort();
    }
    if (status != CDS_OK) { printf("compressed data is corrupt"); fflush(stdout); }   // pg_lzcompress.c:138-139
}

// ---- consumer casts ----------------------------------------------------------------------------

extern "C" int cds_onehot_argmax_i16(const void *d_onehot, uint64_t n_rows, uint32_t classes, void *d_labels, void *stream)
{
    if (n_rows == 0) return CDS_OK;
    if (!d_onehot || !d_labels || classes == 0) return CDS_ERR_ARG;
    uint64_t blocks = (n_rows + 7) / 8;
    const uint64_t max_blocks = (uint64_t)current_sm_count() * 16;
    if (blocks > max_blocks) blocks = max_blocks;
    k_onehot_argmax_i16<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream>>>(static_cast<const int16_t *>(d_onehot), n_rows, classes,
                                                                           static_cast<long long *>(d_labels));
    CU(cudaGetLastError());
    LAUNCHED(1);
    return CDS_OK;
}

// Tile of whole pixels for k_assemble_tma: tp divides hw, tp*ch*4 is a multiple of 16 and at most
// kAsmMaxTileBytes, tp even when two pixels share a bf16 store.  0: the staged kernel does not apply.
static uint32_t assemble_tile(const void *images, uint32_t hw, uint32_t ch, bool bf16)
{
    if ((reinterpret_cast<uintptr_t>(images) & 15u) || (((uint64_t)hw * ch * 4u) & 15u)) return 0;
    if (bf16 && (hw & 1u)) return 0;
    uint32_t best = 0;
    for (uint32_t tp = 1; tp <= hw && (uint64_t)tp * ch * 4u <= kAsmMaxTileBytes; tp++) {
        if (hw % tp) continue;
        if (((uint64_t)tp * ch * 4u) & 15u) continue;
        if (bf16 && (tp & 1u)) continue;
        best = tp;
    }
    return (best && (uint64_t)best * ch * 4u >= 4096u) ? best : 0;
}

template <typename TOut>
static int launch_assemble_tma(const void *d_images, uint64_t n_samples, uint32_t hw, uint32_t ch, uint32_t tp,
                               const void *d_labels, const void *d_indices, uint32_t batch, void *d_out, void *d_out_labels,
                               cudaStream_t st)
{
    const size_t smem = (size_t)kAsmStages * tp * ch * 4u + 2 * kAsmStages * sizeof(unsigned long long) + 16;
    CU(cudaFuncSetAttribute(k_assemble_tma<TOut>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int sms = current_sm_count();
    const uint64_t total = (uint64_t)batch * (hw / tp);
    const uint32_t grid = (uint32_t)(total < (uint64_t)sms ? total : (uint64_t)sms);
    k_assemble_tma<TOut><<<grid, kAsmThreads, smem, st>>>(
        static_cast<const float *>(d_images), n_samples, hw, ch, tp, static_cast<const long long *>(d_labels),
        static_cast<const long long *>(d_indices), batch, static_cast<TOut *>(d_out), static_cast<long long *>(d_out_labels));
    CU(cudaGetLastError());
    LAUNCHED(1);
    return CDS_OK;
}

extern "C" int cds_nhwc_to_nchw_f32(const void *d_src, uint64_t n, uint32_t hw, uint32_t ch, void *d_dst, void *stream)
{
    if (n == 0) return CDS_OK;
    if (!d_src || !d_dst || hw == 0 || ch == 0) return CDS_ERR_ARG;
    if (n <= 0xFFFFFFFFull && ch > 1) {
        const uint32_t tp = assemble_tile(d_src, hw, ch, false);
        if (tp) return launch_assemble_tma<float>(d_src, n, hw, ch, tp, nullptr, nullptr, (uint32_t)n, d_dst, nullptr, (cudaStream_t)stream);
    }
    uint64_t total = n * hw * ch, blocks = (total + 255) / 256;
    const uint64_t max_blocks = (uint64_t)current_sm_count() * 32;
    if (blocks > max_blocks) blocks = max_blocks;
    k_nhwc_to_nchw_f32<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream>>>(static_cast<const float *>(d_src), n, hw, ch,
                                                                          static_cast<float *>(d_dst));
    CU(cudaGetLastError());
    LAUNCHED(1);
    return CDS_OK;
}

extern "C" int cds_f32_to_bf16(const void *d_src, uint64_t n_elems, void *d_dst, void *stream)
{
    if (n_elems == 0) return CDS_OK;
    if (!d_src || !d_dst) return CDS_ERR_ARG;
    uint64_t blocks = (n_elems / 4 + 255) / 256 + 1;
    const uint64_t max_blocks = (uint64_t)current_sm_count() * 32;
    if (blocks > max_blocks) blocks = max_blocks;
    k_f32_to_bf16<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream>>>(static_cast<const float *>(d_src), n_elems,
                                                                     static_cast<__nv_bfloat16 *>(d_dst));
    CU(cudaGetLastError());
    LAUNCHED(1);
    return CDS_OK;
}

extern "C" int cds_batch_assemble(const void *d_images, uint64_t n_samples, uint32_t hw, uint32_t ch, const void *d_labels,
                                  const void *d_indices, uint32_t batch, void *d_out, int out_bf16, void *d_out_labels,
                                  void *stream)
{
    if (batch == 0) return CDS_OK;
    if (!d_images || !d_indices || !d_out || n_samples == 0 || hw == 0 || ch == 0) return CDS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (ch > 1) {
        const uint32_t tp = assemble_tile(d_images, hw, ch, out_bf16 != 0);
        if (tp) {
            if (out_bf16)
                return launch_assemble_tma<__nv_bfloat16>(d_images, n_samples, hw, ch, tp, d_labels, d_indices, batch, d_out, d_out_labels, st);
            return launch_assemble_tma<float>(d_images, n_samples, hw, ch, tp, d_labels, d_indices, batch, d_out, d_out_labels, st);
        }
    }
    uint64_t total = (uint64_t)batch * hw * ch, blocks = (total + 255) / 256;
    const uint64_t max_blocks = (uint64_t)current_sm_count() * 32;
    if (blocks > max_blocks) blocks = max_blocks;
    if (out_bf16)
        k_batch_assemble<__nv_bfloat16><<<(uint32_t)blocks, 256, 0, st>>>(
            static_cast<const float *>(d_images), n_samples, hw, ch, static_cast<const long long *>(d_labels),
            static_cast<const long long *>(d_indices), batch, static_cast<__nv_bfloat16 *>(d_out),
            static_cast<long long *>(d_out_labels));
    else
        k_batch_assemble<float><<<(uint32_t)blocks, 256, 0, st>>>(
            static_cast<const float *>(d_images), n_samples, hw, ch, static_cast<const long long *>(d_labels),
            static_cast<const long long *>(d_indices), batch, static_cast<float *>(d_out),
            static_cast<long long *>(d_out_labels));
    CU(cudaGetLastError());
    LAUNCHED(1);
    return CDS_OK;
}

// ---- checksum (bench.py's correctness gate) and the model hopper's state copy ------------------------

extern "C" int cds_checksum64(const void *d_buf, uint64_t n_bytes, uint64_t *d_out, void *stream)
{
    if (!d_out || (!d_buf && n_bytes)) return CDS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    CU(cudaMemsetAsync(d_out, 0, sizeof(uint64_t), st));
    if (n_bytes == 0) return CDS_OK;
    uint64_t blocks = (n_bytes / 8 + 255) / 256 + 1;
    const uint64_t max_blocks = (uint64_t)current_sm_count() * 8;
    if (blocks > max_blocks) blocks = max_blocks;
    k_checksum64<<<(uint32_t)blocks, 256, 0, st>>>(static_cast<const uint8_t *>(d_buf), n_bytes, reinterpret_cast<unsigned long long *>(d_out));
    CU(cudaGetLastError());
    LAUNCHED(1);
    return CDS_OK;
}

extern "C" int cds_peer_copy(const void *d_src, void *d_dst, uint64_t n_bytes, void *stream)
{
    if (n_bytes == 0) return CDS_OK;
    if (!d_src || !d_dst) return CDS_ERR_ARG;
    uint64_t blocks = (n_bytes / 16 + 511) / 512 + 1;
    const uint64_t max_blocks = (uint64_t)current_sm_count() * 4;
    if (blocks > max_blocks) blocks = max_blocks;
    k_peer_copy<<<(uint32_t)blocks, 512, 0, (cudaStream_t)stream>>>(static_cast<const uint8_t *>(d_src), static_cast<uint8_t *>(d_dst), n_bytes);
    CU(cudaGetLastError());
    LAUNCHED(1);
    return CDS_OK;
}

// ---- text ETL -----------------------------------------------------------------

extern "C" int cds_text_etl(int device, const void *d_text, uint64_t n_bytes, uint32_t values_per_row, uint32_t label_digits,
                            void *d_features, void *d_labels, void *d_ids, uint64_t cap_rows, uint64_t *n_rows,
                            uint64_t *err_pos, void *stream)
{
    if (!n_rows || values_per_row == 0 || (n_bytes && (!d_text || !d_features || !d_labels))) return CDS_ERR_ARG;
    *n_rows = 0;
    if (err_pos) *err_pos = 0;
    if (n_bytes == 0) return CDS_OK;
    if (cds_device_count() == 0) return CDS_ERR_NO_DEVICE;
    cudaStream_t st = (cudaStream_t)stream;
    ON_DEVICE(device);
    const uint64_t n_tiles64 = (n_bytes + kTextTile - 1) / kTextTile;
    if (n_tiles64 > 0x7FFFFFFFull) return CDS_ERR_CAPACITY;
    const uint32_t n_tiles = (uint32_t)n_tiles64, scan_tiles = n_tiles / kScanMinTile + 2;
    const uint32_t per_row = values_per_row + 2u;
    // work area: [TextState | 2 tickets | 2 x status] zeroed, then counts, then scan values (grow-only, see WorkArea)
    if (device < 0 || device >= 64) return CDS_ERR_ARG;
    const size_t zero_bytes = align256(256 + 4 * (8 + 2 * (size_t)scan_tiles));
    const size_t count_bytes = align256(sizeof(uint32_t) * 2 * ((size_t)n_tiles + 1)), val_bytes = align256(sizeof(unsigned long long) * 4 * (size_t)scan_tiles);
    std::lock_guard<std::mutex> lock(g_work_mu[kWorkText]);
    uint8_t *zero = nullptr;
    cudaError_t e = work_reserve(device, kWorkText, zero_bytes + count_bytes + val_bytes, &zero);
    uint32_t *counts = reinterpret_cast<uint32_t *>(zero + zero_bytes);
    unsigned long long *vals = reinterpret_cast<unsigned long long *>(zero + zero_bytes + count_bytes);
    TextState h;
    memset(&h, 0, sizeof(h));
    if (e == cudaSuccess) e = cudaMemsetAsync(zero, 0, zero_bytes, st);
    if (e == cudaSuccess) {
        TextState *state = reinterpret_cast<TextState *>(zero);
        uint32_t *w = reinterpret_cast<uint32_t *>(zero + 256);
        ScanWs ws0{w, w + 8, vals, vals + scan_tiles}, ws1{w + 1, w + 8 + scan_tiles, vals + 2 * scan_tiles, vals + 3 * scan_tiles};
        uint32_t *tile_tokens = counts, *tile_lines = counts + n_tiles + 1;
        const uint8_t *text = static_cast<const uint8_t *>(d_text);
        k_text_count<<<n_tiles, kTextThreads, 0, st>>>(text, n_bytes, tile_tokens, tile_lines);
        uint32_t sg = n_tiles / (kScanThreads * 4) + 1;
        const uint32_t sg_max = (uint32_t)sm_count(device) * 4u;
        if (sg > sg_max) sg = sg_max;
        k_text_scan<<<dim3(sg, 2), kScanThreads, 0, st>>>(tile_tokens, tile_lines, n_tiles, ws0, ws1, state, n_bytes, text);
        k_text_parse<<<n_tiles, kTextThreads, 0, st>>>(text, n_bytes, tile_tokens, tile_lines, per_row, label_digits,
                                                       static_cast<float *>(d_features), static_cast<long long *>(d_labels),
                                                       static_cast<long long *>(d_ids), cap_rows, state);
        e = cudaGetLastError();
        LAUNCHED(3);
        if (e == cudaSuccess) e = cudaMemcpyAsync(&h, state, sizeof(h), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    }
    if (e != cudaSuccess) return cuda_fail(e, "cds_text_etl");
    *n_rows = h.n_lines;
    if (h.err_inv) {
        const unsigned long long key = ~h.err_inv;
        if (err_pos) *err_pos = (((key >> 28) & 0x1FFFFFFFFull) << 4) | ((key >> 8) & 0xFu);
        return (int)(key & 0xFFu);
    }
    if (h.n_tokens != h.n_lines * per_row) return CDS_ERR_SHAPE;            // the last line has no newline to check it
    return CDS_OK;
}
