/* quorum_b200 — C ABI of the B200 (sm_100a) thinking-tag filter + concatenate path.
 *
 * The reference (andrewginns/quorum) has no FFI: the hot path is reached through two module-level Python
 * names and two inline `str.join` expressions in src/quorum/oai_proxy.py.  This header is what a binding for
 * that path binds instead; every entry point cites the reference code it replaces.  All pointers are plain
 * pointers + sizes (no torch types).  Unless stated otherwise every data pointer is a DEVICE pointer, all
 * offsets are int64 byte offsets into the flat buffer they index, and the call is asynchronous on `stream`
 * (a cudaStream_t passed as void*).  The library allocates no device memory: scratch is caller-owned, its size
 * comes from the *_scratch_bytes query.
 *
 * Return value: 0 on success, a negative QB200_E* code otherwise; qb200_last_error() gives the message of the
 * last failing call on the calling thread.  There is no CPU fallback: without a CUDA device every compute
 * entry point fails with QB200_ECUDA.
 *
 * Text must be (generalised) UTF-8: the automata look only at ASCII bytes and the multi-byte forms listed in
 * DESIGN.md; every other byte passes through untouched.
 */
#ifndef QUORUM_B200_H
#define QUORUM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QB200_OK 0
#define QB200_EINVAL (-1)   /* bad argument (NULL, negative count, misaligned base, capacity too small) */
#define QB200_ETAGS (-2)    /* unsupported tag table (see qb200_tags_create) */
#define QB200_ECUDA (-3)    /* CUDA runtime error (no device, launch failure) */
#define QB200_ESCRATCH (-4) /* scratch buffer too small */

#define QB200_MAX_TAGS 16    /* distinct (case-folded) names */
#define QB200_MAX_TAG_LEN 22 /* bytes per name */
#define QB200_HELD_MAX 23    /* longest partial tag a stream can carry, in bytes */

/* flags */
#define QB200_HIDE 1u        /* strip_concat: apply the filter + .strip()  (hide_final_think, oai_proxy.py:134-139) */
#define QB200_STREAM_JOIN 2u /* strip_concat: streaming-final join, drop inputs that are empty BEFORE the filter (oai_proxy.py:760-764) */
#define QB200_FLUSH 4u       /* stream_feed: apply ThinkingTagFilter.flush() after the last chunk (oai_proxy.py:739, 352-371) */
#define QB200_SKIP_EMPTY 16u /* sse_frame: items with empty content produce no event (`if safe_text`, oai_proxy.py:625-628) */
#define QB200_CHAIN 8u       /* this call is one slice of a larger batch, see qb200_stream_feed_batch / qb200_strip_concat_batch */

typedef struct qb200_tags qb200_tags; /* opaque, host-side */

/* Per-stream carried state of ThinkingTagFilter (oai_proxy.py:275-276: self.buffer, self.thinking_depth).
 * `held` is the part of the reference's buffer that can still matter: at depth 0 a holdable partial open tag
 * (oai_proxy.py:311-323), at depth>0 a partial open/close tag at the very end (oai_proxy.py:332-334).
 * `tail_len` lets a host mirror `self.buffer` exactly: number of bytes fed by this call after the last tag it
 * consumed, or -1 when it consumed none. */
typedef struct {
    int32_t depth;
    int32_t tail_len;
    uint8_t held_len;
    uint8_t held[QB200_HELD_MAX];
} qb200_stream_state; /* 32 bytes */

/* Thread-local message of the last failing call. */
const char *qb200_last_error(void);

/* Build the tag table.  Replaces the pattern construction of strip_thinking_tags (oai_proxy.py:137-138) and
 * ThinkingTagFilter.__init__ (oai_proxy.py:270-276): names are case-folded and de-duplicated; an empty list means
 * the pattern "<()>" with hold-back disabled, exactly as the reference behaves.
 * `names` (HOST pointer): concatenated name bytes; `name_off` (HOST, n_tags+1 entries).
 * Supported names: printable ASCII without regex metacharacters and without '<' '>' '/', at most QB200_MAX_TAG_LEN
 * bytes, at most QB200_MAX_TAGS distinct names, at most 128 backreference classes in total (2^(number of 'i' and
 * 's' letters) per name).  Anything else is rejected with QB200_ETAGS rather than approximated. */
int qb200_tags_create(const uint8_t *names, const int32_t *name_off, int32_t n_tags, qb200_tags **out);
void qb200_tags_destroy(qb200_tags *tags);

/* strip_thinking_tags over n_resp responses fused with the concatenate join over n_req requests.
 *   replaces  strip_thinking_tags                       oai_proxy.py:120-139
 *             non-stream concatenate                    oai_proxy.py:1194-1199 + 1286 (1279, 1284)
 *             streaming-final concatenate               oai_proxy.py:759-764 + 841 (834, 839)   [QB200_STREAM_JOIN;
 *                                                       the caller passes "\n"+separator as `sep`]
 * Response r is in_bytes[in_off[r] .. in_off[r+1]).  Request q joins responses req_off[q] .. req_off[q+1]-1 in
 * order with `sep` between consecutive INCLUDED responses (all of them, or with QB200_STREAM_JOIN those whose
 * input length is > 0).  req_off == NULL means one response per request (n_req must equal n_resp): plain batched
 * strip_thinking_tags.  Without QB200_HIDE the bytes are copied unchanged (hide flag False, oai_proxy.py:134-135).
 * Outputs: request q's text is out_bytes[out_off[q] .. out_off[q+1]) (out_off has n_req+1 entries, out_off[0]=0);
 * resp_out_len (optional, n_resp entries) receives each response's filtered length (0 for excluded ones).
 * out_capacity must be >= in_off[n_resp] + sep_len * n_resp.   in_bytes and out_bytes must be 16-byte aligned.
 * (out_capacity also sizes the scratch tables, see the *_scratch_bytes query.  A caller that knows its output is shorter and
 * states a capacity below the input size still gets exact results as long as out_bytes holds the output: the units that
 * lie beyond the stated bound are walked from their bytes, the slow way, instead of from the token records.)
 * QB200_CHAIN feeds one large batch slice by slice (to overlap host<->device copies with compute): in_bytes and
 * out_bytes stay those of the whole batch, `in_off` points at the slice's first response (entries still absolute byte
 * offsets), `req_off` holds the slice's requests as response indices relative to that first response, `out_off` points at
 * the slice's first request and out_off[0] already holds where the slice's output starts (the previous slice's call wrote
 * it; the first slice is called without the flag).  Slices are issued in order on one stream. */
size_t qb200_strip_concat_scratch_bytes(int64_t n_resp, int64_t n_req, int64_t out_capacity);
int qb200_strip_concat_batch(const qb200_tags *tags, const uint8_t *in_bytes, const int64_t *in_off, int64_t n_resp,
                             const int64_t *req_off, int64_t n_req, const uint8_t *sep, int32_t sep_len,
                             uint32_t flags, uint8_t *out_bytes, int64_t out_capacity, int64_t *out_off,
                             int64_t *resp_out_len, void *scratch, size_t scratch_bytes, void *stream);

/* ThinkingTagFilter.feed over many streams x many chunks, state carried in and out.
 *   replaces  ThinkingTagFilter.feed                    oai_proxy.py:278-350
 *             ThinkingTagFilter.flush                   oai_proxy.py:352-371   [QB200_FLUSH]
 *             the per-backend filter objects            oai_proxy.py:543-545, 618-624, 739-741
 * Chunk c is chunk_bytes[chunk_off[c] .. chunk_off[c+1]); stream s owns chunks stream_off[s] .. stream_off[s+1]-1,
 * fed in order (stream_off[0] = 0, stream_off[n_streams] = n_chunks; a stream's chunks are contiguous).
 * state_in == NULL starts every stream fresh (depth 0, nothing held); state_out may alias state_in or be NULL.
 * Outputs: the return value of the feed of chunk c is out_bytes[out_chunk_off[c] .. out_chunk_off[c+1])
 * (n_chunks+1 entries).  out_capacity must be >= chunk_off[n_chunks] + QB200_HELD_MAX * n_streams.
 * QB200_CHAIN lets a host feed one large batch slice by slice (e.g. to overlap host<->device copies with compute): all
 * arrays stay those of the WHOLE batch, `stream_off` points at the slice's first stream (its entries are still absolute
 * chunk indices), `n_streams` is the slice's stream count, state pointers point at the slice's first stream, and the
 * slice's output starts at out_chunk_off[stream_off[0]], which the previous slice's call wrote (the first slice is called
 * without the flag).  Slices must be issued in order on one stream; only the slice's bytes and chunk offsets need to be
 * resident, plus chunk_off[n_chunks] (the total, which bounds every read). */
size_t qb200_stream_feed_scratch_bytes(int64_t n_chunks, int64_t n_streams, int64_t out_capacity);
int qb200_stream_feed_batch(const qb200_tags *tags, const uint8_t *chunk_bytes, const int64_t *chunk_off,
                            int64_t n_chunks, const int64_t *stream_off, int64_t n_streams,
                            const qb200_stream_state *state_in, qb200_stream_state *state_out, uint32_t flags,
                            uint8_t *out_bytes, int64_t out_capacity, int64_t *out_chunk_off, void *scratch,
                            size_t scratch_bytes, void *stream);

/* Same computation as qb200_stream_feed_batch for callers that cannot provide scratch memory.  No counterpart in the
 * reference (it is a property of this library's deployment, not of quorum's algorithm).
 * Meant for hosts that cannot allocate at call time: a process whose device memory is entirely pre-reserved by its
 * own pools, a binding that owns only its input/output buffers, or a loop where cudaMalloc's device-wide
 * synchronisation is not acceptable.  The scratch is a fixed array inside the library (about 1 MiB of device memory on
 * every GPU that loads it, held until the process exits), so:
 *   - at most QB200_NOALLOC_MAX_STREAMS streams per call (larger batches: QB200_EINVAL);
 *   - all calls on one device share that array; calls that may overlap on the GPU race on it: serialise them;
 *   - the first call on each device looks up the array's address (do it once, with that device current, before any
 *     stream capture begins).
 * The call itself performs no cudaMalloc / cudaFree and calls no allocating library routine.  Arguments and outputs are
 * those of qb200_stream_feed_batch without `scratch` / `scratch_bytes`. */
#define QB200_NOALLOC_MAX_STREAMS 65536
int qb200_stream_feed_batch_noalloc(const qb200_tags *tags, const uint8_t *chunk_bytes, const int64_t *chunk_off,
                                    int64_t n_chunks, const int64_t *stream_off, int64_t n_streams,
                                    const qb200_stream_state *state_in, qb200_stream_state *state_out, uint32_t flags,
                                    uint8_t *out_bytes, int64_t out_capacity, int64_t *out_chunk_off, void *stream);

/* ThinkingTagFilter.flush for n streams without feeding (oai_proxy.py:352-371): drops the held partial, keeps depth. */
int qb200_stream_flush_batch(qb200_stream_state *state, int64_t n_streams, void *stream);

/* SSE output framing of filtered deltas (SURVEY.md section 8, row f1 - the component right after the filter).
 *   replaces  f"data: {json.dumps(stream_event)}\n\n".encode()      oai_proxy.py:629-652 (per-backend events)
 *             f"data: {json.dumps(final_event)}\n\n".encode()       oai_proxy.py:847-862 (final event)
 * Everything in the event except delta.content is constant per backend, so the host renders, once, one PREFIX per
 * backend (`data: {"id": "chatcmpl-parallel-<i>", ..., "delta": {"content": "`) and one SUFFIX (`"}, "finish_reason": null}]}\n\n`)
 * with json.dumps; item k becomes prefix[item_prefix[k]] + json_escape(content_k) + suffix, where json_escape is CPython's
 * ensure_ascii string encoder (short forms for " \ \n \r \t \b \f, \u00xx for the other C0 controls and 0x7F, \uXXXX for
 * every non-ASCII code point, surrogate pairs above U+FFFF, lower-case hex).  With QB200_SKIP_EMPTY an item whose content
 * is empty produces no bytes.  item_prefix may be NULL (all items use prefix 0).  All pointers are device pointers;
 * event k is out_bytes[out_off[k] .. out_off[k+1]).  out_capacity must be >= 6 * content bytes + n_items * (longest
 * prefix + suffix_len). */
size_t qb200_sse_frame_scratch_bytes(int64_t n_items);
int qb200_sse_frame_batch(const uint8_t *in_bytes, const int64_t *in_off, int64_t n_items, const int32_t *item_prefix,
                          const uint8_t *prefix_bytes, const int64_t *prefix_off, int32_t n_prefix, const uint8_t *suffix,
                          int32_t suffix_len, uint32_t flags, uint8_t *out_bytes, int64_t out_capacity, int64_t *out_off,
                          void *scratch, size_t scratch_bytes, void *stream);

/* SSE input parsing, replaces what src/quorum/oai_proxy.py:578-617 does with every HTTP chunk
 * of a backend's stream before the filter: strict UTF-8 decode of the chunk (a chunk that does not decode has no events),
 * .strip().split("\n\n"), startswith("data: "), [6:].strip(), "[DONE]", json.loads, choices[0].delta.content.
 * Two calls, because the number of events is only known after the first:
 *   qb200_sse_split_batch  -> ev_first[c] = index of chunk c's first event, ev_first[n_http] = number of events
 *   (the host reads that one value, sizes out_off / ev_kind, and calls)
 *   qb200_sse_parse_batch  -> one entry per event, in order: ev_kind[k] = 0 no delta, 1 delta (its JSON-unescaped content
 *   is out_bytes[out_off[k] .. out_off[k+1]); lone surrogates in their 3-byte form), 2 the host must decide (anything
 *   that is not certainly valid and well-typed under CPython's json module: there the reference's behaviour is Python's
 *   exception semantics); *n_host (device) = number of kind-2 events.  Events of kind 0 and 2 have empty content, so
 *   (out_bytes, out_off) can go to qb200_stream_feed_batch as (chunk_bytes, chunk_off) directly when *n_host is 0.
 * out_capacity >= http_off[n_http] (unescaping never grows a string).  All pointers are device pointers. */
size_t qb200_sse_split_scratch_bytes(int64_t n_http);
int qb200_sse_split_batch(const uint8_t *http_bytes, const int64_t *http_off, int64_t n_http, int64_t *ev_first, void *scratch,
                          size_t scratch_bytes, void *stream);
size_t qb200_sse_parse_scratch_bytes(int64_t n_events);
int qb200_sse_parse_batch(const uint8_t *http_bytes, const int64_t *http_off, int64_t n_http, const int64_t *ev_first,
                          int64_t n_events, uint8_t *out_bytes, int64_t out_capacity, int64_t *out_off, uint8_t *ev_kind,
                          int64_t *n_host, void *scratch, size_t scratch_bytes, void *stream);

/* Aggregate-strategy prompt assembly, replaces the loop + join + template substitution of
 * src/quorum/oai_proxy.py:406-423 over a batch of requests:
 *   prompt[q] = head[q] + join_{i}(label[i] + response[req_off[q] + i], sep) + tail
 * where the host renders head[q] = query_format.format(query=...) + (template text before "{responses}"), tail = the
 * template text after it, label[i] = source_label_format.format(backend_name=f"LLM{i+1}") (n_labels = 0 when
 * include_source_names is false; positions i >= n_labels get no label).  The responses are the already filtered contents
 * (device buffer, e.g. the output of qb200_strip_concat_batch with one response per request).  head_off may be NULL (no
 * heads).  All pointers are device pointers.  out_capacity >= sum of the prompt lengths; out_off has n_req + 1 entries. */
size_t qb200_assemble_scratch_bytes(int64_t n_resp, int64_t n_req);
int qb200_assemble_prompts_batch(const uint8_t *resp_bytes, const int64_t *resp_off, int64_t n_resp, const int64_t *req_off,
                                 int64_t n_req, const uint8_t *head_bytes, const int64_t *head_off,
                                 const uint8_t *label_bytes, const int32_t *label_off, int32_t n_labels, const uint8_t *sep,
                                 int32_t sep_len, const uint8_t *tail, int32_t tail_len, uint8_t *out_bytes,
                                 int64_t out_capacity, int64_t *out_off, void *scratch, size_t scratch_bytes, void *stream);

/* Usage sums of the combined response (row f4), replaces src/quorum/oai_proxy.py:1298-1311: usage is [n_resp][3]
 * (prompt_tokens, completion_tokens, total_tokens), out is [n_req][3], request q covers responses
 * [req_off[q], req_off[q+1]).  Device pointers. */
int qb200_usage_sums_batch(const int64_t *usage, const int64_t *req_off, int64_t n_req, int64_t *out, void *stream);

/* 32-bit lengths <-> 64-bit offsets on the device.  No counterpart in the reference (it has no buffers): the offset arrays of
 * the two batch entry points are int64, which for 64-byte chunks is 11 % of what a host uploads and 20 % of what it reads
 * back; a host can ship int32 lengths over PCIe instead and convert here.
 *   qb200_offsets_from_lengths32: off[i + 1] = off[0] + len[0] + .. + len[i] for i < n; off[0] must already hold the start
 *     (0 for a whole batch; for a slice of a larger batch the previous slice's call left it there).
 *   qb200_lengths32_from_offsets: len[i] = off[i + 1] - off[i].
 * Device pointers. */
size_t qb200_offsets_scratch_bytes(int64_t n);
int qb200_offsets_from_lengths32(const int32_t *len, int64_t n, int64_t *off, void *scratch, size_t scratch_bytes, void *stream);
int qb200_lengths32_from_offsets(const int64_t *off, int64_t n, int32_t *len, void *stream);

/* Number of kernels this library has launched in the calling process (bench.py's gpu_launches). */
int64_t qb200_launch_count(void);

/* Per-kernel split of the calls made between the two functions, without a profiler attached: while enabled, the library
 * records a CUDA event on the launching stream after each of its kernel launches.  qb200_profile_end synchronises on the
 * last event and returns the number of launches seen (at most max_entries are written): names[i] is a static string
 * (the kernel), ms[i] the time from the previous event to this one.  Measurement aid for bench.py; process-wide;
 * launches from several host threads are recorded safely (one list, mutex), begin / end themselves are for one thread at a
 * time; no counterpart in the reference. */
int qb200_profile_begin(void *stream);
int32_t qb200_profile_end(const char **names, float *ms, int32_t max_entries);

#ifdef __cplusplus
}
#endif
#endif
