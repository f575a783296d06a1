/*
 * cds_da.h — C ABI of libcds_da.so: the B200 (sm_100a) Direct-Access reader for Cerebro-DS.
 *
 * Every entry point is plain `extern "C"`: pointers, sizes, an `int` status.  No torch types.
 * Device pointers are raw CUDA device addresses (e.g. tensor.data_ptr()); `stream` is a
 * cudaStream_t passed as void* (NULL = the legacy default stream).
 *
 * Each declaration names the reference code it replaces; paths are relative to
 * /root/reference/cerebro_gpdb/ ("ppr" = pg_page_reader.py).  The reference's own FFI for this
 * path is ctypes (ppr:47-57: wrap_function(lib, 'pglz_decompress', None, [c_char_p, c_char_p]));
 * INTEGRATION.md shows the ctypes stubs a maintainer of the reference would add.
 */
#ifndef CDS_DA_H
#define CDS_DA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CDS_BLCKSZ 32768u               /* ppr:34  BLOCK_SIZE                */
#define CDS_TOAST_MAX_CHUNK_SIZE 8140u  /* ppr:44  TOAST_MAX_CHUNK_SIZE      */
#define CDS_ROW_WORDS 16u               /* u32 words per exported heap row   */

/* Status codes.  1..14 are data errors and map one-to-one onto what the reference raises
 * (the Python shim re-raises the same exception type and message); >=100 are infrastructure. */
enum cds_status {
    CDS_OK = 0,
    CDS_ERR_PARTIAL_PAGE = 1,   /* Exception("FATAL")                               ppr:375-377 */
    CDS_ERR_ITEMID_ALIGN = 2,   /* Exception("Item identifier length do not match") ppr:262-265 */
    CDS_ERR_TOAST_SPECIAL = 3,  /* AssertionError("THERE SHALL NOT BE INDICES")     ppr:388-389 */
    CDS_ERR_TUPLE = 4,          /* struct.error: tuple/chunk does not fit its page  ppr:273,404  */
    CDS_ERR_RAWSIZE = 5,        /* AssertionError('raw size does not match!')       ppr:539      */
    CDS_ERR_CHUNK_COUNT = 6,    /* AssertionError (numchunks)                       ppr:572      */
    CDS_ERR_CHUNK_EXTENDED = 7, /* AssertionError (chunk header is extended)        ppr:580      */
    CDS_ERR_MAX_CHUNK = 8,      /* Exception("max chunk size error")                ppr:584-586  */
    CDS_ERR_CHUNK_SIZE = 9,     /* Exception("unexpected chunk size")               ppr:587-591  */
    CDS_ERR_FINAL_SIZE = 10,    /* AssertionError("final size does not match")      ppr:595-596  */
    CDS_ERR_PGLZ_CORRUPT = 11,  /* Exception("compressed data is corrupt")          ppr:229-230  */
    CDS_ERR_SHAPE = 12,         /* ValueError (array size vs. catalog shape)        ppr:167,173  */
    CDS_ERR_CATALOG_MISS = 13,  /* IndexError (.iloc[0] on an empty frame)          ppr:563-564  */
    CDS_ERR_MISSING_VALUE = 14, /* a pointer whose value has no chunks (reference: None entry)  */
    CDS_ERR_TEXT_NUMBER = 15,   /* text ETL: NaN/Infinity/malformed literal (literal_eval raises too) or a
                                   literal outside the exactly convertible range (see cds_text_etl)  */
    CDS_ERR_CUDA = 100,         /* a CUDA runtime call failed: see cds_last_error()             */
    CDS_ERR_ARG = 101,          /* bad argument                                                  */
    CDS_ERR_CAPACITY = 102,     /* rows / chunks / scratch / output larger than the capacity    */
    CDS_ERR_NO_DEVICE = 103     /* no CUDA device: this library has no CPU path                  */
};

typedef struct cds_segment cds_segment; /* one per (GPU, segment being decoded); not thread-safe */

typedef struct cds_summary {
    uint64_t n_rows;        /* heap tuples (packed buffers) seen                            */
    uint64_t n_positions;   /* entries of the result dict                                   */
    uint64_t ind_bytes;     /* decoded bytes of all independent_var values                  */
    uint64_t dep_bytes;     /* decoded bytes of all dependent_var values                    */
    uint64_t n_chunks;      /* TOAST tuples walked                                          */
    uint64_t n_inline;      /* rows whose dependent_var is inline-compressed                */
    uint64_t scratch_bytes; /* staging bytes used by compressed external values             */
    uint64_t err_where;     /* page / row / chunk index the first error refers to           */
    uint32_t status;        /* first data error in the reference's processing order, or 0   */
    uint32_t err_stage;     /* 0 heap scan, 1 TOAST scan, 2 inline rows, 3 external, 4 plan */
    uint64_t n_wide_values; /* compressed values large enough for the intra-value parallel decoder  */
} cds_summary;

/* ---- library ------------------------------------------------------------------------------ */
int cds_version(void);
uint32_t cds_summary_size(void);                /* sizeof(cds_summary) as the library was built: bindings check their mirror */
const char *cds_status_str(int status);
const char *cds_last_error(void);               /* text of the last CUDA failure in this thread */
int cds_device_count(void);

/* Pinned host memory for page staging (async H2D needs it).  Free with cds_host_free. */
int cds_host_alloc(void **ptr, uint64_t bytes);
int cds_host_free(void *ptr);

/* ---- segment handle ------------------------------------------------------------------------
 * All device memory the decode needs is allocated here, once: capacities are in pages.
 * max_chunks = 0 -> 6 per TOAST page + 4096.  scratch_bytes = ~0 -> as many bytes as the TOAST
 * capacity (always enough); 0 -> none (segments with compressed external values are refused).
 * stage_pages != 0 also allocates device page buffers so that cds_decode_host can be used.
 * With scratch, the work area of the intra-value parallel pglz decoder is allocated too: about
 * 0.18 bytes per scratch byte (sub-segment entries, segment tables, block flags). */
int cds_segment_create(cds_segment **out, int device, uint64_t max_heap_pages, uint64_t max_toast_pages,
                       uint64_t max_chunks, uint64_t scratch_bytes, int stage_pages);
int cds_segment_destroy(cds_segment *seg);

/* Output format of independent_var for the following cds_materialize / cds_decode_* calls on this handle:
 * 0 (default) = the stored bytes (float32 for these tables; a bit copy, as ppr:165-182),
 * 1 = bfloat16, converted from the stored float32 with round-to-nearest-even inside the gather and the
 *     pglz decoder (no float32 copy is written); capacities and cds_summary.ind_bytes stay in float32
 *     bytes, the buffer needs ind_bytes / 2.  Not in the reference: an optional training-buffer dtype.
 * 2 = the stored bytes are uint8 pixel components (HWC, `channels` interleaved) and leave as float32
 *     lut[channel][x] — the normalisation the reference applied before loading,
 *     ((x / 255.0) - mean[c]) / std[c] (preprocessing/imagenet/generate_h5_file.py:77-81), made exact by
 *     evaluating it once per table entry in double on the host (cds_set_normalize_lut first).  The buffer
 *     needs 4 * ind_bytes; values must start on a channel boundary (whole samples do). */
int cds_set_output_format(cds_segment *seg, int independent_var_format);
/* The [channels][256] float32 table for format 2 (channels <= 4), copied to the device. */
int cds_set_normalize_lut(cds_segment *seg, const float *h_lut, uint32_t channels);

/* ---- the path, step by step (each enqueues on `stream` and returns; no host sync) ---------- */

/* table_page_read (ppr:451-494): page header (ppr:253-269), ItemId (ppr:285-305), tuple header
 * (ppr:272-276) and MADlib packed-row parse (ppr:328-351) of `n_pages` heap pages on the device. */
int cds_heap_scan(cds_segment *seg, const void *d_heap_pages, uint64_t n_pages, void *stream);

/* Alternative to cds_heap_scan for callers that already hold the rows (toast_page_read takes a
 * df_toast argument, ppr:615): upload `n_rows` rows in the cds_export_rows layout; inline datums
 * are addressed relative to `d_inline_blob`. */
int cds_set_rows(cds_segment *seg, const uint32_t *h_rows, uint64_t n_rows, const void *d_inline_blob,
                 void *stream);

/* read_toast_all_bytes (ppr:497-524) over generator_page(toast=True) (ppr:385-422): physical
 * tuple walk of `n_pages` TOAST pages, chunk directory, chunk_id -> pointer matching. */
int cds_toast_scan(cds_segment *seg, const void *d_toast_pages, uint64_t n_pages, void *stream);

/* Bookkeeping half of detoast (ppr:527-612): result order (inline rows in table order, then
 * external values in first-seen order), decoded sizes, output offsets.  *_sample_bytes may be 0;
 * when both are set every buffer must hold equally many samples in both values
 * (run_da_cerebro_standalone.py:43). */
int cds_plan(cds_segment *seg, uint32_t ind_sample_bytes, uint32_t dep_sample_bytes, void *stream);

/* Data half of detoast: chunk checks (ppr:570-596), chunk gather, pglz decode (ppr:191-231 ==
 * pg_lzcompress.c:49-143) and the typed reinterpretation (ppr:165-182, a byte copy) straight
 * into caller-owned device buffers: values back to back in result order. */
int cds_materialize(cds_segment *seg, void *d_out_ind, uint64_t ind_cap_bytes, void *d_out_dep,
                    uint64_t dep_cap_bytes, void *stream);

/* Wait for everything enqueued for this segment on `stream` and fetch counts + first error.
 * Returns summary->status (a data error) or an infrastructure error. */
int cds_finish(cds_segment *seg, void *stream, cds_summary *summary);

/* ---- the path, whole ------------------------------------------------------------------------ */

/* da.input_fn (da.py:55-58) with pages already in HBM: heap_scan + toast_scan + plan +
 * materialize.  summary == NULL: enqueue only (call cds_finish later). */
int cds_decode_device(cds_segment *seg, const void *d_heap_pages, uint64_t n_heap_pages,
                      const void *d_toast_pages, uint64_t n_toast_pages,
                      uint32_t ind_sample_bytes, uint32_t dep_sample_bytes,
                      void *d_out_ind, uint64_t ind_cap_bytes, void *d_out_dep, uint64_t dep_cap_bytes,
                      void *stream, cds_summary *summary);

/* Same with HOST page buffers (pinned for async copies).  Needs a handle created with stage_pages != 0.
 * The decode is overlapped with the copy (the reference reads and parses page by page, ppr:358-379): the heap goes
 * first, the TOAST relation follows in ranges of pages on the library's own copy stream, and every range is indexed,
 * walked, gathered and its completed values decompressed on `stream` as soon as it has arrived.  That needs the plan
 * before the TOAST relation is there; it is taken from the heap alone, assuming that external values lie in the TOAST
 * relation in table order and that a value decodes to va_rawsize - 4 bytes — true of every table a loader writes.  The
 * complete chunk directory is checked against that plan at the end; if it does not bear it out, or any error was
 * met, cds_finish decodes the segment again the classic way from the pages (all on the device by then), so results,
 * status codes and their order are those of cds_decode_device in every case.  As with the other entry points, the
 * outputs of an enqueue-only call (summary == NULL) are final once cds_finish has returned.
 * Not streamed: relations of three ranges or fewer (the copy is too short to hide anything), a `stream` that is
 * being captured, a call with timing enabled. */
int cds_decode_host(cds_segment *seg, const void *h_heap_pages, uint64_t n_heap_pages,
                    const void *h_toast_pages, uint64_t n_toast_pages,
                    uint32_t ind_sample_bytes, uint32_t dep_sample_bytes,
                    void *d_out_ind, uint64_t ind_cap_bytes, void *d_out_dep, uint64_t dep_cap_bytes,
                    void *stream, cds_summary *summary);

/* Pages per range of the streamed cds_decode_host (default 2048 = 64 MiB; 0 = copy everything, then decode).
 * range_pages | 0x80000000 streams every relation of more than one range (tests). */
int cds_set_host_streaming(cds_segment *seg, uint32_t range_pages);
/* cds_decode_host calls of this process that were streamed, and how many of those cds_finish had to decode again. */
int cds_stream_counts(unsigned long long *streamed, unsigned long long *decoded_again);

/* ---- measurement hooks (bench.py) ------------------------------------------------------------ */

/* Kernels launched by this library in this process so far. */
unsigned long long cds_launch_count(void);
/* Record CUDA-event pairs on the launching stream inside the next `max_samples` cds_materialize / cds_decode_*
 * calls (0 disables).  Stages: 0 = the gather kernel; 1 = the pglz stage (classification, wide decoder, warp
 * decoder); 2, 3 = the block kernels of the wide decoder in front of and behind its tail kernels (k_wide_blocks<2> and
 * k_wide_finish with a store of resolved blocks, the two parses k_wide_blocks<1> and <3> without; when its kernels are
 * enqueued with the decode).
 * cds_timing_read_stage waits for the samples of one stage and returns their durations in milliseconds;
 * cds_timing_read does the same for stage 0 and then rearms (call it last). */
int cds_timing_enable(cds_segment *seg, uint32_t max_samples);
int cds_timing_read(cds_segment *seg, float *h_ms, uint32_t cap, uint32_t *n);
int cds_timing_read_stage(cds_segment *seg, uint32_t stage, float *h_ms, uint32_t cap, uint32_t *n);

/* ---- results for the Python mirror of the reader interface (synchronous, small) ------------ */

/* df_data / df_toast material (ppr:477-492).  Row = 16 u32: dist_key, buffer_id,
 * independent_var {header bytes, va_rawsize, va_extsize, va_valueid, va_toastrelid},
 * dependent_var {same five}, dep_external, inline_off_lo, inline_off_hi, inline_len
 * (byte offset/length of tupdata[24:] inside the heap buffer, ppr:344). */
int cds_export_rows(cds_segment *seg, uint32_t *h_rows, uint64_t cap_rows, uint64_t *n_rows, void *stream);

/* Result layout after cds_plan: order[p] = row at dict position p; per ROW byte offset/length of
 * its values inside the two output buffers.  Arrays hold n_positions / n_rows entries. */
int cds_export_layout(cds_segment *seg, uint64_t *h_order, uint64_t *h_ind_off, uint64_t *h_dep_off,
                      uint64_t *h_ind_len, uint64_t *h_dep_len, uint64_t cap_rows, void *stream);

/* Chunk directory after cds_toast_scan, in the reference's yield order (ppr:412-422):
 * chunk_id, chunk_seq, byte offset of the payload inside the TOAST buffer, payload length. */
int cds_export_chunks(cds_segment *seg, uint32_t *h_chunk_id, uint32_t *h_chunk_seq, uint64_t *h_payload_off,
                      uint32_t *h_payload_len, uint64_t cap_chunks, uint64_t *n_chunks, void *stream);

/* ---- pglz on its own ------------------------------------------------------------------------ */

/* Batch pglz_decompress (ppr:191-231, pg_lzcompress.c:49-143) on device buffers: datum i is the
 * varlena (8-byte header + stream) at d_src + src_off[i], src_off[i+1]-src_off[i] bytes long,
 * decoded to d_dst + dst_off[i] (dst_off[i+1]-dst_off[i] bytes of room).  h_status[i] receives
 * CDS_OK or CDS_ERR_PGLZ_CORRUPT.  Offsets are host arrays of n+1 entries.  Synchronous.
 * Its work memory (and that of pglz_decompress and cds_text_etl) is one grow-only device allocation per
 * device, kept between calls: repeated calls of the same size never reach cudaMalloc.  Calls are
 * serialised by a lock. */
int cds_pglz_decompress_batch(int device, const void *d_src, const uint64_t *h_src_off, void *d_dst,
                              const uint64_t *h_dst_off, uint64_t n, int32_t *h_status, void *stream);

/* Intra-value parallel pglz (the paper's packed buffers, one value = 3210 images = 483 MB,
 * load_imagenet.py:27-31,118-135).  pglz_decompress (ppr:191-231, pg_lzcompress.c:49-143) is serial inside a
 * value; values whose decoded size is at least the threshold can be decoded by the whole GPU instead of by one
 * warp (csrc/pglz_wide.cuh), same bytes, same corrupt/ok verdict.  Whether they are is decided on the device,
 * per decode, from the sizes in the job list: few huge values go wide, many medium ones (a thousand 7 MB
 * Criteo buffers) keep one warp each, which is faster then.  Process-wide settings: threshold in decoded bytes
 * (default 1 MiB; 0 switches the wide decoder off; values under 1024 bytes never take it);
 * skip_cost_model != 0 sends every value at or above the threshold through it (tests, measurements). */
int cds_set_wide_pglz_threshold(uint32_t decoded_bytes, int skip_cost_model);
/* The wide decoder parses every 48 KB output block once when it can keep the resolved block (2 bytes per decoded byte:
 * the byte, or which byte of the previous block's last 4095 it copies) until the blocks' tails are known, and twice when
 * it cannot.  cds_pglz_decompress_batch / pglz_decompress take that store from their cached work area; a handle gets
 * one from cds_finish after the first decode that sent values to the wide decoder, sized for that decode plus an eighth
 * (replaced by a larger one when a later decode needs it, freed by cds_segment_destroy), so the decodes that follow on
 * the handle take the one-parse path.  Process-wide limit in bytes for any one store (default: none beyond "at most half
 * of the free device memory"; 0: no store, always two parses).  Same bytes and verdicts either way.  No counterpart in
 * the reference. */
int cds_set_wide_store_limit(uint64_t bytes);
/* Per handle: how the pglz stage of a decode is launched.  Which values go wide is always decided on the device
 * for THIS decode (one small classification kernel after the gather); the modes differ in whether the 15 kernels
 * of the wide decoder are enqueued along with it.
 *   2 (default) = they are enqueued when the handle's previous summary saw values go wide, or the handle has not
 *       produced a summary yet (a performance hint only).  When they were not enqueued and the classification does take values, cds_finish runs them before it
 *       returns — one extra round trip on the first such segment, never a slow decode and never a wrong one; the
 *       outputs of an enqueue-only decode (summary == NULL) are complete once cds_finish has returned.
 *   0 = warp decoder only.
 *   1 = always enqueued, and every value at or above the threshold goes wide (tests, measurements).
 *   3 = always enqueued, cost model on (profiling a steady state). */
int cds_set_wide_pglz(cds_segment *seg, int mode);
/* Frees the cached work memory of cds_pglz_decompress_batch / pglz_decompress / cds_text_etl on `device`. */
int cds_trim_work_memory(int device);
/* Measurement hook: values decoded by the wide decoder so far in this process (counted when a decode's
 * summary is read / a batch call returns). */
unsigned long long cds_wide_decoded_count(void);

/* No counterpart in the reference.  cds_pglz_decompress_batch for callers that are not allowed to
 * allocate device memory while they decode — a loader thread beside a framework that has claimed the
 * whole GPU memory pool, a latency-sensitive ingest loop in which an allocation could stall the device,
 * a process with a hard memory budget.  Same bytes, same per-datum status, always one warp per datum (the
 * intra-value parallel decoder needs a work area that grows with the data, so it is never used here: a value of
 * hundreds of megabytes takes seconds on its warp — send those through a segment handle).  Every pointer is a DEVICE pointer the
 * caller already owns (offsets are uint64[n+1] on the device; d_status int32[n]; d_first_corrupt, if
 * not NULL, receives the index of the first corrupt datum or -1) and the work area is an array built
 * into the library (one copy, ~128 KB, on every GPU that loads it, held until the process exits).
 * Limits and rules:
 *   - at most CDS_STATIC_MAX_DATUMS datums per call; more is refused with CDS_ERR_CAPACITY and a
 *     message in cds_last_error();
 *   - the call only enqueues work on `stream` (memset + kernels) and never calls cudaMalloc/cudaFree (the
 *     CUDA runtime itself may allocate when a device is first used or the module is first loaded: make the
 *     first call per device at start-up);
 *   - the first call on a device looks up the work area's address (cudaGetSymbolAddress, with that
 *     device current): make that first call outside any stream capture.  The address is cached for the life
 *     of the process: do not call again after cudaDeviceReset;
 *   - all calls on a device share the one work area: two calls whose kernels can overlap on the GPU
 *     (different streams, no dependency) race with each other — serialise them. */
#define CDS_STATIC_MAX_DATUMS 4096u
int cds_pglz_decompress_batch_static(int device, const void *d_src, const uint64_t *d_src_off, void *d_dst,
                                     const uint64_t *d_dst_off, uint64_t n, int32_t *d_status,
                                     int64_t *d_first_corrupt, void *stream);

/* The reference's own symbol and call shape (pg_lzcompress.c:49-50; ctypes binding at ppr:47-49,
 * destination from ctypes.create_string_buffer(rawsize), ppr:247-248): HOST pointers, no status
 * (the C reference printf()s on corruption; this prints the same text to stdout).  Runs on the
 * current CUDA device — there is no CPU decoder in this library. */
void pglz_decompress(const void *source, char *dest);

/* ---- consumer-side casts (a14) -------------------------------------------------------------- */

/* np.argmax(dependent_var[idx]) (run_pytorchddp_da.py:55): first maximum of each int16 row. */
int cds_onehot_argmax_i16(const void *d_onehot, uint64_t n_rows, uint32_t classes, void *d_labels_i64,
                          void *stream);
/* np.rollaxis(image, 2, 0) per sample + collate (run_pytorchddp_da.py:56-58): NHWC -> NCHW. */
int cds_nhwc_to_nchw_f32(const void *d_src, uint64_t n, uint32_t hw, uint32_t channels, void *d_dst,
                         void *stream);
/* Optional training-buffer dtype (not in the reference; round-to-nearest-even). */
int cds_f32_to_bf16(const void *d_src, uint64_t n_elems, void *d_dst, void *stream);

/* DADataset.__getitem__ + DataLoader collate (run_pytorchddp_da.py:51-62, run_pytorchddp.py:397-425)
 * as one launch: sample i = indices[b] mod n_samples (the train-mode epoch padding of
 * run_pytorchddp_da.py:52-53), out[b] = NCHW view of images[i] (f32, or bf16 when out_bf16 != 0),
 * out_labels[b] = labels[i].  d_labels / d_out_labels may be NULL.  indices: int64 on the device. */
int cds_batch_assemble(const void *d_images, uint64_t n_samples, uint32_t hw, uint32_t channels,
                       const void *d_labels, const void *d_indices, uint32_t batch, void *d_out, int out_bf16,
                       void *d_out_labels, void *stream);

/* ---- checksum and model-state copy ------------------------------------------------------------ */

/* Order-sensitive 64-bit checksum of a device buffer into *d_out (device memory): the sum over its 64-bit
 * little-endian words w_i (tail zero-padded) of w_i * (2 i + 1) * 0x9E3779B97F4A7C15 modulo 2^64.  bench.py compares it
 * with the same sum over the arrays the segment generator packed (the reference's result compared as a whole,
 * run_da_cerebro_standalone.py:38-56 consumes exactly these bytes). */
int cds_checksum64(const void *d_buf, uint64_t n_bytes, uint64_t *d_out, void *stream);

/* The model hopper's transfer (the reference moves the flat float32 state of
 * madlib_keras_wrapper.py:119-131 between workers as files, ctq.py:330-333,404-405): copies n_bytes from a buffer
 * on the current device to d_dst, which may be another GPU's buffer mapped into this process (symmetric memory /
 * CUDA IPC).  One kernel of plain 16-byte loads and stores over NVLink; enqueues on `stream`. */
int cds_peer_copy(const void *d_src, void *d_dst, uint64_t n_bytes, void *stream);

/* ---- text ETL ("next" row f-4) ---------------------------------------------------------------- */

/* The parser of the filesystem ETL (etl_imagenet.py:30-37 `parser`, etl_criteo.py:28-37) for gpfdist
 * TEXT unloads (unload_imagenet.sql:1-14): lines `id|{v,v,...}|label` in device memory -> features
 * float32 [rows, values_per_row] (nested braces are flattened in file order, as np.asarray of the nested
 * list), labels int64 [rows] (label_digits == 0: int(label), etl_imagenet.py:36) or int64 [rows,
 * label_digits] (label is a digit string, one int per character, etl_criteo.py:35-36), ids int64 [rows]
 * (may be NULL).  Numbers follow the reference's double rounding: float(token) exactly (decimal ->
 * binary64, correctly rounded: Eisel-Lemire with Clinger's path as back-up) then binary64 -> float32.
 * Accepted: at most 19 significant digits and a decimal exponent (after shifting the digits into the
 * mantissa) in [-80, 60] — every float4/float8 text value of the unloads; anything else, NaN and
 * Infinity return CDS_ERR_TEXT_NUMBER with the byte position in *err_pos.  A line whose token count is
 * not values_per_row + 2 returns CDS_ERR_SHAPE.  Synchronous; *n_rows = lines parsed. */
int cds_text_etl(int device, const void *d_text, uint64_t n_bytes, uint32_t values_per_row, uint32_t label_digits,
                 void *d_features, void *d_labels, void *d_ids, uint64_t cap_rows, uint64_t *n_rows,
                 uint64_t *err_pos, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* CDS_DA_H */
