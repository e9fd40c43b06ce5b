// Aggregate-strategy prompt assembly and the usage sums of the combined response (f4).
//
// oai_proxy.py:406-423 builds, per request,
//     query_format.format(query=user_query) + prompt_template.replace("{responses}", separator.join(label_i + response_i))
// The host renders the two per-request constants (head = the formatted query + the template text before "{responses}",
// tail = the template text after it) and the label table; the device does the part that scales with the responses:
// lengths -> prefix sum -> every response, label and separator copied once to its final place.
#pragma once
#include "qb_common.cuh"

namespace qb {

// thread per request: length of its prompt; relative position of each response's segment ([sep] label response)
__global__ void assemble_layout_kernel(const int64_t* __restrict__ resp_off, const int64_t* __restrict__ req_off, int64_t n_req,
                                       const int64_t* __restrict__ head_off, const int32_t* __restrict__ label_off, int32_t n_labels,
                                       int32_t sep_len, int32_t tail_len, int64_t* __restrict__ rel, int32_t* __restrict__ idx,
                                       int64_t* __restrict__ reqlen) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_req) return;
    const int64_t r0 = req_off[q], r1 = req_off[q + 1];
    int64_t o = head_off ? head_off[q + 1] - head_off[q] : 0;
    for (int64_t r = r0; r < r1; r++) {
        const int i = (int)(r - r0);
        rel[r] = o;
        idx[r] = i;
        if (i > 0) o += sep_len;
        if (i < n_labels) o += label_off[i + 1] - label_off[i];
        o += resp_off[r + 1] - resp_off[r];
    }
    reqlen[q] = o + tail_len;
}

// warp per request: head and tail to their place
__global__ void __launch_bounds__(256)
assemble_frame_kernel(const int64_t* __restrict__ req_off, int64_t n_req, const uint8_t* __restrict__ head_bytes,
                      const int64_t* __restrict__ head_off, const uint8_t* __restrict__ tail, int32_t tail_len,
                      const int64_t* __restrict__ out_off, uint8_t* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t q = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (q >= n_req) return;
    if (head_off) {
        const int64_t a = head_off[q], n = head_off[q + 1] - a;
        for (int64_t i = lane; i < n; i += 32) out[out_off[q] + i] = head_bytes[a + i];
    }
    const int64_t e = out_off[q + 1];
    for (int i = lane; i < tail_len; i += 32) out[e - tail_len + i] = tail[i];
}

// warp per response: [separator] [label] response
__global__ void __launch_bounds__(256)
assemble_copy_kernel(const uint8_t* __restrict__ resp_bytes, const int64_t* __restrict__ resp_off, int64_t n_resp,
                     const int64_t* __restrict__ rel, const int32_t* __restrict__ idx, const int64_t* __restrict__ req_of,
                     const uint8_t* __restrict__ label_bytes, const int32_t* __restrict__ label_off, int32_t n_labels,
                     const uint8_t* __restrict__ sep, int32_t sep_len, const int64_t* __restrict__ out_off,
                     uint8_t* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t total = resp_off[n_resp];
    for (int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < n_resp; r += nw) {
        int64_t d = out_off[req_of[r]] + rel[r];
        const int i = idx[r];
        if (i > 0) {
            for (int k = lane; k < sep_len; k += 32) out[d + k] = sep[k];
            d += sep_len;
        }
        if (i < n_labels) {
            const int a = label_off[i], n = label_off[i + 1] - a;
            for (int k = lane; k < n; k += 32) out[d + k] = label_bytes[a + k];
            d += n;
        }
        const int64_t a = resp_off[r], n = resp_off[r + 1] - a;
        if (n > 0) warp_copy<false>(out + d, resp_bytes + a, n, resp_bytes + total, lane);
    }
}

// thread per request: which request a response belongs to (the copy kernel is response-parallel)
__global__ void assemble_owner_kernel(const int64_t* __restrict__ req_off, int64_t n_req, int64_t* __restrict__ req_of) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_req) return;
    for (int64_t r = req_off[q]; r < req_off[q + 1]; r++) req_of[r] = q;
}

// f4 (oai_proxy.py:1298-1311): per request, the sums of prompt_tokens / completion_tokens / total_tokens over its responses.
// usage is [n_resp][3], out is [n_req][3].
__global__ void usage_sums_kernel(const int64_t* __restrict__ usage, const int64_t* __restrict__ req_off, int64_t n_req,
                                  int64_t* __restrict__ out) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_req) return;
    int64_t a = 0, b = 0, c = 0;
    for (int64_t r = req_off[q]; r < req_off[q + 1]; r++) {
        a += usage[3 * r];
        b += usage[3 * r + 1];
        c += usage[3 * r + 2];
    }
    out[3 * q] = a;
    out[3 * q + 1] = b;
    out[3 * q + 2] = c;
}

}  // namespace qb
