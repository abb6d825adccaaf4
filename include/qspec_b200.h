/*
 * qspec_b200.h -- C ABI of the B200-native QSpec dual-mode int4 linear path.
 *
 * One shared library (qspec_b200/_C/libqspec_b200.so), plain pointers + sizes +
 * a CUDA stream handle, `int` status returns. No torch types cross this boundary.
 * Every pointer is a DEVICE pointer unless the name ends in `_host`. All tensors are
 * contiguous row-major. Every launch goes to the `stream` argument (a cudaStream_t
 * cast to void*; NULL = legacy default stream). Nothing here synchronises.
 *
 * Each entry point names the reference interface it replaces (paths relative to the
 * hku-netexplo-lab/QSpec tree). The reference-side binding a maintainer would add is
 * shown in INTEGRATION.md.
 *
 * Nibble format everywhere (reference: third-party/QuaRot/quarot/functional/quantization.py:41-74):
 *   signed int4 in [-8,7], two's complement; byte j of a row = (q[2j] & 0xF) | (q[2j+1] << 4).
 */
#ifndef QSPEC_B200_H_
#define QSPEC_B200_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes ------------------------------------------------------------------- */
#define QSPEC_OK                 0
#define QSPEC_ERR_INVALID_ARG    1   /* null pointer, non-positive size                    */
#define QSPEC_ERR_UNSUPPORTED    2   /* shape/dtype outside what the kernels implement      */
#define QSPEC_ERR_CUDA           3   /* a CUDA runtime/driver call failed (see last_error)  */
#define QSPEC_ERR_WORKSPACE      4   /* workspace pointer null or too small                 */

/* 16-bit float element type of activations / outputs / scales. */
#define QSPEC_DTYPE_F16   0
#define QSPEC_DTYPE_BF16  1

/* ABI version, bumped on any signature change. */
int  qspec_abi_version(void);
/* Human-readable text for the last non-OK status returned on the calling thread. */
const char* qspec_last_error(void);
/* Number of kernels this library has launched in the calling process (bench's gpu_launches). */
uint64_t qspec_launch_count(void);
/* SM count of the current device (cached), 0 when no device is usable. */
int  qspec_sm_count(void);
/* Debug only: when non-null, the GEMM kernels of CTA 0 and the last CTA record %globaltimer
 * timestamps / cycle counters into this device buffer (8 regions of 4096 uint64, one per launch, round robin).
 * Only the profiling build (python -m qspec_b200.build --prof) records anything. */
void qspec_debug_set_gemm_trace(void* device_buffer);

/* ---- (A) norm (+ int4 quant) --------------------------------------------------------
 * Replaces qserve_backend.layernorm_ops.rms_norm_general_fuse_sum_i4
 *   (third-party/kernels/csrc/layernorm.cpp:32-38; host layernorm_kernels.cu:890-924;
 *    kernel :569-720) as called by quarot_nn/normalization.py:34-65.
 * Per row (fp32 math, fixed reduction tree shared with oracle/):
 *   mean = sum(x)/K; var = sum((x-mean)^2)/K; r = 1/sqrt(var+eps); y = (x-mean)*r;
 *   yh = f16(y); amax = max(|yh|, f16(1e-6)); q = rni(clamp(y*(7/amax), -8, 7)) packed;
 *   scale[row] = f16(amax/7); input_sum[row] = f16(sum(yh)).
 * x [rows,K] f16; out_q [rows,K/2] int8; scale,input_sum [rows] f16 (input_sum may be NULL).
 * K % 16 == 0, K <= 16384.
 */
int qspec_norm_quant_i4(const void* x, int8_t* out_q, void* scale, void* input_sum,
                        int rows, int K, float eps, void* stream);

/* Replaces layernorm_ops.rms_norm_general_fuse_sum_fp16 (layernorm.cpp:40-42; host
 * layernorm_kernels.cu:927-957; same kernel, branch :679-682), caller normalization.py:67-82.
 * out[row,i] = f16((x-mean)*r). x,out [rows,K] of `dtype` (F16 is the reference's). */
int qspec_norm_f16(const void* x, void* out, int rows, int K, float eps, int dtype, void* stream);

/* Fused residual add + norm (quarot_llama.py:380,389 `hidden = residual + hidden` followed by
 * quarot_nn/normalization.py:34-82): x <- T(x + delta) is written back in place, then normalised (and, in the
 * draft mode, quantised) exactly as qspec_norm_quant_i4 / qspec_norm_f16 would on the updated x. Bit-identical to
 * qspec_add followed by the unfused norm; saves one launch and one round trip of the hidden state. */
int qspec_add_norm_quant_i4(void* x, const void* delta, int8_t* out_q, void* scale, void* input_sum,
                            int rows, int K, float eps, void* stream);
int qspec_add_norm_f16(void* x, const void* delta, void* out, int rows, int K, float eps, int dtype,
                       void* stream);

/* ---- deferred GEMM reduction: the k-split sum of o_proj / down_proj done by the norm that consumes it ----
 * qspec_gemm_w4a4_split / qspec_gemm_w4a16_split run the same GEMM as qspec_gemm_w4a4 / qspec_gemm_w4a16 but leave
 * the RAW partial accumulators of the CTAs that share an output tile: partials [split][M][N] 32-bit words
 * (W4A4: int32, 256 x the true partial sum; W4A16: fp32), split = qspec_gemm_split_factor(w4a16, M, N, K)
 * (0: shape not supported in this mode, use the plain call). qspec_split_add_norm_* then do
 *   delta = T(epilogue(sum over slices)), x += delta, norm (+ quantiser)
 * with exactly the arithmetic and summation order of the in-GEMM reduction: for W4A4 (exact int32 sums) x and the norm
 * outputs are bit-identical to qspec_gemm_* followed by qspec_add_norm_*; for W4A16 they are when both calls split k the
 * same way, and equal up to fp32 summation order otherwise. What it saves: the GEMM's cross-CTA reduction
 * (partial store, counter atomic, reload: three serialized L2 round trips at the end of every launch).
 * Reference semantics: quarot_nn/linear.py:34-56 (Linear4bit.forward) + quarot_llama.py:380,389 (residual add)
 * + layernorm_kernels.cu:569-720. */
int qspec_gemm_split_factor(int w4a16, int M, int N, int K);
/* Introspection of the GEMM scheduler (tests, tuning scripts; no reference counterpart, no launch): mode 0 = W4A4, 1 = W4A16
 * f16, 2 = W4A16 bf16, 3 = W4A8. out[0..7] = {token tile, CTA pairs?, grid (CTAs or CTA pairs), cluster size, deferred
 * split, ring stages, token tiles, (pair-)tiles along N}. */
int qspec_debug_gemm_plan(int mode, int M, int N, int K, int deferred, int* out);
int qspec_gemm_w4a4_split(const int8_t* a, const int8_t* w, void* partials, size_t partial_bytes, int M, int N, int K,
                          void* workspace, size_t workspace_bytes, void* stream);
int qspec_gemm_w4a16_split(const void* x, const int8_t* w, void* partials, size_t partial_bytes, int M, int N, int K,
                           int dtype, void* workspace, size_t workspace_bytes, void* stream);
int qspec_split_add_norm_quant_i4(void* x, const void* partials, int split, int w4a16, const void* sA, const void* sW,
                                  const void* bias, int8_t* out_q, void* scale, void* input_sum, int rows, int K,
                                  float eps, void* stream);
int qspec_split_add_norm_f16(void* x, const void* partials, int split, int w4a16, const void* sA, const void* sW,
                             const void* bias, void* out, int rows, int K, float eps, int dtype, void* stream);

/* ---- int8 container for the draft path's quantised activations ("_c8") ------------------------------------------
 * Same reference lines as the functions they shadow (the quantised activation is QuaRot's PackedQuantizedTensor,
 * quarot_nn/quantization.py:10-19; the GEMM is Linear4bit.forward_w4a4, quarot_nn/linear.py:63-84): identical arithmetic and
 * identical int4 VALUES, but written one per byte ([rows, K] int8 in [-8, 7]) instead of two per byte, so that the GEMM takes
 * them through its int8-activation instantiation (qspec_gemm_w4a8 / qspec_gemm_w4a8_split: TMA delivers the operand as it
 * lies) and does not widen the activation tile again for every weight tile. Outputs are bit-identical to the packed route:
 * acc = 16 * sum(a * w) >> 4 instead of 256 * sum(a * w) >> 8. Worth it from ~64 token rows (12.8 vs 14.5 us for the 8B qkv
 * GEMM, profiles/r02y); the packed format stays the interchange format of the plugin API. In qspec_gemm_split_factor and
 * qspec_split_add_norm_* the kind argument (`w4a16`) takes 2 for partials produced by qspec_gemm_w4a8_split. */
int qspec_split_add_norm_quant_i4_c8(void* x, const void* partials, int split, int w4a16, const void* sA, const void* sW,
                                     const void* bias, int8_t* out_q, void* scale, void* input_sum, int rows, int K,
                                     float eps, void* stream);
int qspec_tp_add_norm_quant_i4_c8(void* x, const void* const* peer_partials, void* const* peer_flags, void* state,
                                  int world, int rank, int8_t* out_q, void* scale, void* input_sum,
                                  int rows, int K, float eps, void* stream);
int qspec_hadamard_heads_c8(const void* x, void* out_f16, int8_t* out_q, void* scale_out,
                            int rows, int H, int D, float scale, float clip_ratio, void* stream);
int qspec_hadamard_k28_c8(const void* x, const void* gate_up, void* out_f16, int8_t* out_q,
                          void* scale_out, int rows, int n, float scale, float clip_ratio, void* stream);
int qspec_gemm_w4a8_split(const int8_t* a, const int8_t* w, void* partials, size_t partial_bytes, int M, int N, int K,
                          void* workspace, size_t workspace_bytes, void* stream);

/* ---- tensor parallel: all-reduce of the row-parallel partial sums fused into the norm -----
 * Replaces tensor_model_parallel_all_reduce in RowParallelLinear.forward (vllm layers/linear.py:1143-1144)
 * plus the residual add and the norm that follow it (quarot_llama.py:380,389 and :374/:387): one launch per
 * rank, no collective call. peer_partials[r] is rank r's [rows, K] partial output of o_proj / down_proj in
 * memory this GPU can address (NVLink peer mapping, own rank included); peer_flags[r] is rank r's array of
 * QSPEC_TP_MAX zero-initialised 32-bit arrival flags; state is 4 zero-initialised 32-bit words of LOCAL memory
 * {epoch, CTAs done, time-outs, reserved}. delta = T(sum_r partial_r), summed in rank order in fp32 and rounded
 * once; then exactly qspec_add_norm_*(x, delta, ...). Every rank must issue the same sequence of tp calls on
 * the same rows, alternating between two partial buffers (a rank may overwrite buffer A only after its next tp
 * call on buffer B was issued). state[2] counts waits that gave up after 30 s (a peer died): check it. */
#define QSPEC_TP_MAX 8
int qspec_tp_add_norm_quant_i4(void* x, const void* const* peer_partials, void* const* peer_flags, void* state,
                               int world, int rank, int8_t* out_q, void* scale, void* input_sum,
                               int rows, int K, float eps, void* stream);
int qspec_tp_add_norm_f16(void* x, const void* const* peer_partials, void* const* peer_flags, void* state,
                          int world, int rank, void* out, int rows, int K, float eps, int dtype, void* stream);

/* Two-step form of the same fused all-reduce for large exchanges (same reference lines; same result bit for bit): row r is
 * summed by rank r % world only, which pushes the rounded row into every peer's peer_reduced[] buffer ([rows, K], storage type,
 * 16-byte aligned) and publishes the epoch in the peer's peer_row_flags[] word of that row (>= rows zero-initialised 32-bit words per
 * rank; the arrival flags peer_flags and `state` are shared with the one-step form). The other ranks read the row from local memory:
 * 2 (world-1)/world x rows x K x 2 bytes of NVLink traffic per rank instead of (world-1) x rows x K x 2. rows <= 1024.
 * phases: 3 = the product (one launch); 1 / 2 = owned rows / other rows as separate launches (single-GPU emulation in tests:
 * every rank's phase 1 before any rank's phase 2). All ranks must use the same form for the same exchange. */
int qspec_tp_rs_add_norm_quant_i4(void* x, const void* const* peer_partials, void* const* peer_flags,
                                  void* const* peer_reduced, void* const* peer_row_flags, void* state, int world, int rank,
                                  int phases, int8_t* out_q, void* scale, void* input_sum, int rows, int K, float eps,
                                  void* stream);
int qspec_tp_rs_add_norm_f16(void* x, const void* const* peer_partials, void* const* peer_flags,
                             void* const* peer_reduced, void* const* peer_row_flags, void* state, int world, int rank,
                             int phases, void* out, int rows, int K, float eps, int dtype, void* stream);

/* Greedy token over a vocabulary-sharded lm_head without a collective call: replaces the all-gather of
 * LogitsProcessor._get_logits (vllm/model_executor/layers/logits_processor.py:84-107: tensor_model_parallel_gather /
 * all_gather of the sharded logits) followed by the sampler's arg max (layers/sampler.py greedy path) for the greedy decode this path runs.
 * local_max / local_idx [rows]: maximum logit (f32) and its LOCAL index (int64) of this rank's shard (qspec_argmax_rows);
 * index_offset = rank * shard size. peer_slots[r]: rank r's table of 2 * QSPEC_TP_MAX * capacity (f32, u32) pairs in memory
 * this GPU can address, zero-initialised; peer_flags[r]: rank r's QSPEC_TP_MAX zero-initialised 32-bit arrival flags (NOT
 * the ones of qspec_tp_add_norm_*); state: 4 zero-initialised words of local memory {epoch, -, time-outs, -}. Every rank
 * pushes its pairs into every table, then takes the best of its own in rank order (ties: lowest rank = lowest global index).
 * phases: 3 = the product (push + combine); 1 / 2 = the two halves as separate launches (single-GPU emulation in tests).
 * out_tokens [rows] int64, identical on every rank. */
int qspec_tp_argmax_combine(const float* local_max, const long long* local_idx, long long index_offset, int rows,
                            void* const* peer_slots, void* const* peer_flags, void* state, int capacity,
                            int world, int rank, int phases, long long* out_tokens, void* stream);

/* ---- per-token symmetric int4 quant ---------------------------------------------------
 * Replaces quarot._CUDA.fuse_sym_quant (third-party/QuaRot/quarot/kernels/bindings.cpp:128-146,
 * kernel quant.cu:102-186), caller quarot_nn/quantization.py:10-19.
 *   amax = max|x| (f16); scale = f16(f16(amax/7) * f16(clip)); q = clamp(rni(f16(x/scale)),-8,7).
 * x [rows,K] f16; scale [rows] f16; out_q [rows,K/2]. K % 16 == 0, K <= 32768.
 */
int qspec_sym_quant_i4(const void* x, void* scale, int8_t* out_q, int rows, int K,
                       float clip_ratio, void* stream);

/* ---- W4A8 mode (off by default in the reference: kwargs["w4a8"]) -----------------------------
 * qspec_norm_quant_i8 replaces layernorm_ops.rms_norm_general_fuse_sum_i8 (third-party/kernels/csrc/layernorm.cpp:44-50, kernel
 * layernorm_kernels.cu:350-470), caller quarot_nn/normalization.py:84-105: the norm of qspec_norm_quant_i4 with
 * s = 127/amax, q = cvt.rni.sat.s8(y*s), one int8 per value: out_q [rows,K]; scale = f16(amax/127).
 * qspec_sym_quant_i8 replaces quarot._CUDA.fuse_sym_quant_i8 (bindings.cpp:148-164, kernel quant.cu:188-245), caller
 * quarot_nn/quantization.py:20-26: scale = f16(f16(amax/127)*f16(clip)); q = clamp(rni(f16(x/scale)),-128,127); out_q [rows,K].
 * qspec_gemm_w4a8 replaces torchao's rowwise_scaled_linear_cutlass_s8s4 as called by Linear4bit.forward_w4a8
 * (quarot_nn/linear.py:86-99): a [M,K] int8, everything else as qspec_gemm_w4a4 (same weights, same epilogue). */
int qspec_norm_quant_i8(const void* x, int8_t* out_q, void* scale, void* input_sum,
                        int rows, int K, float eps, void* stream);
int qspec_sym_quant_i8(const void* x, void* scale, int8_t* out_q, int rows, int K,
                       float clip_ratio, void* stream);
size_t qspec_gemm_w4a8_workspace_bytes(int M, int N, int K);
int qspec_gemm_w4a8(const int8_t* a, const void* sA, const int8_t* w, const void* sW,
                    const void* bias, void* out, int M, int N, int K, int dtype,
                    void* workspace, size_t workspace_bytes, void* stream);
int qspec_gemm_w4a8_acc_i32(const int8_t* a, const int8_t* w, int32_t* acc, int M, int N, int K,
                            void* workspace, size_t workspace_bytes, void* stream);

/* ---- online Hadamard -------------------------------------------------------------------
 * Replaces fast_hadamard_transform_cuda.faster_fast_hadamard_transform(x, scale, out)
 * (third-party/fast-hadamard-transform/csrc/fast_hadamard_transform.cpp:69-110; kernel
 * fast_hadamard_transform_cuda.cu:124-198): Sylvester FHT along the last dim, fp32
 * butterflies in ascending stride, out = T(v*scale). n = power of two, 8 <= n <= 2048.
 */
int qspec_fht(const void* x, void* out, int rows, int n, float scale, int dtype, void* stream);

/* o_proj online Hadamard across heads, with the two transposes folded in.
 * Replaces quarot_llama.py:230-234 (transpose -> OnlineHadamard(K=1) -> transpose -> contiguous)
 * and, when out_q != NULL, also the Quantizer that follows (quarot_llama.py:236-238).
 *   x [rows, H, D] f16 (attention output); y[r,h',d] = f16( sum_h Had_H[h',h] x[r,h,d] * scale ).
 * out_f16 (may be NULL) [rows, H*D]; out_q (may be NULL) [rows, H*D/2] + scale_out [rows] f16 get the
 * sym-quant (clip_ratio) of y. H power of two in [2,64]; D % 16 == 0; H*D <= 16384.
 */
int qspec_hadamard_heads(const void* x, void* out_f16, int8_t* out_q, void* scale_out,
                         int rows, int H, int D, float scale, float clip_ratio, void* stream);

/* down_proj online Hadamard (had28 (x) H_n), optionally fused with SiLU(gate)*up in front and
 * the Quantizer behind. Replaces quarot_nn/hadamard.py:44-49 ->
 * quarot/functional/hadamard.py:109-113 (FHT_n with scale, f16 round, then had28 @ [rows,28,n],
 * f32 accumulate in ascending k, f16 out) and quarot_llama.py:284-294.
 *   x   [rows, 28*n] f16  -- used when gate_up == NULL
 *   gate_up [rows, 2*28*n] f16, layout [up ; gate] (quarot_llama.py:279): a = f16(f16(silu(gate))*up)
 * out_f16 / out_q / scale_out as above (either may be NULL, not both). n power of two in [64,1024].
 */
int qspec_hadamard_k28(const void* x, const void* gate_up, void* out_f16, int8_t* out_q,
                       void* scale_out, int rows, int n, float scale, float clip_ratio,
                       void* stream);

/* ---- W4A4 GEMM (draft mode) ---------------------------------------------------------------
 * Replaces torch.ops.torchao.rowwise_scaled_linear_cutlass_s4s4_unified(xq, x_scale, wq,
 * w_scale, bias, result) (third-party/ao/torchao/ops.py:28-30,600-636; kernel
 * .../rowwise_scaled_linear_cutlass_unified.cuh:240-488), caller quarot_nn/linear.py:67-84.
 *   acc[m,n] = sum_k a[m,k]*w[n,k] (exact int32);
 *   out[m,n] = T( ((float(acc)*float(sA[m]))*float(sW[n])) + float(bias[n]) )
 * a [M,K/2] packed int4; sA [M] T; w [N,K/2] packed int4; sW [N] T; bias [N] T or NULL;
 * out [M,N] T (T = dtype). K % 128 == 0, N % 8 == 0, M >= 1.
 * workspace: qspec_gemm_w4a4_workspace_bytes(M,N,K) bytes, zero-filled once by the caller;
 * the kernel leaves it zeroed on exit (self-cleaning split-K accumulators + tile counters).
 */
size_t qspec_gemm_w4a4_workspace_bytes(int M, int N, int K);
int qspec_gemm_w4a4(const int8_t* a, const void* sA, const int8_t* w, const void* sW,
                    const void* bias, void* out, int M, int N, int K, int dtype,
                    void* workspace, size_t workspace_bytes, void* stream);
/* Debug/parity entry: raw int32 accumulators acc [M,N] (same tcgen05 main loop, no scaling). */
int qspec_gemm_w4a4_acc_i32(const int8_t* a, const int8_t* w, int32_t* acc, int M, int N, int K,
                            void* workspace, size_t workspace_bytes, void* stream);

/* ---- W4A16 GEMM (verify mode) -----------------------------------------------------------------
 * Replaces the BitBLAS Matmul op + the per-call `weight ^ 0x88` pass
 * (quarot_nn/linear.py:102-124,156-211): out = x @ (int4(w) * sW[:,None])^T (+bias),
 * two's-complement nibbles decoded directly, fp32 accumulation.
 * x [M,K] T; w [N,K/2]; sW [N] T; bias [N] T or NULL; out [M,N] T. K % 128 == 0, N % 16 == 0.
 * workspace: qspec_gemm_w4a16_workspace_bytes(M,N,K) bytes zero-filled once (fp32 split-K
 * partials + tile counters, self-cleaning).
 */
size_t qspec_gemm_w4a16_workspace_bytes(int M, int N, int K);
int qspec_gemm_w4a16(const void* x, const int8_t* w, const void* sW, const void* bias, void* out,
                     int M, int N, int K, int dtype, void* workspace, size_t workspace_bytes,
                     void* stream);

/* ---- neighbours of the path used by the decode loop -------------------------------
 * out = a + b elementwise (residual add, quarot_llama.py:380,389). n elements of dtype. */
int qspec_add(const void* a, const void* b, void* out, size_t n, int dtype, void* stream);
/* out[r,i] = f16(f16(silu(gate[r,i])) * up[r,i]) with gate_up [rows, 2*inter] = [up ; gate]
 * (quarot_llama.py:279-284). */
int qspec_silu_mul(const void* gate_up, void* out, int rows, int inter, int dtype, void* stream);

/* RoPE (neox, cos_sin_cache [max_pos, D] = [cos | sin] as vllm ops.rotary_embedding,
 * quarot_llama.py:207-211) applied in place to the q and k slices of qkv [T, (Hq+2Hkv)*D] f16,
 * plus append of rotated k and of v to the contiguous KV cache [B, S, Hkv, D] f16 at
 * (seq_ids[t], positions[t]) (the reference's reshape_and_cache_flash, attention/backends/flash_attn.py:709-718). */
int qspec_rope_kvcache(void* qkv, const void* cos_sin_cache, const int* positions, const int* seq_ids,
                       void* cache_k, void* cache_v, int T, int Hq, int Hkv, int D, int S, void* stream);
/* Causal decode attention over the cache (stands in for the reference's flash-attn call,
 * quarot_llama.py:221): token t attends keys 0..positions[t] of sequence seq_ids[t].
 * out [T, Hq, D] f16. D == 128. */
int qspec_decode_attention(const void* qkv, const void* cache_k, const void* cache_v,
                           const int* positions, const int* seq_ids, void* out, int T, int Hq,
                           int Hkv, int D, int S, float sm_scale, void* stream);

/* qspec_decode_attention for passes with many sequences / several tokens per sequence (verify passes of large batches,
 * BASELINE configs[2]/[3]): one CTA per (sequence, kv head) reads that head's keys and values once for ALL tokens of the
 * sequence in the pass and all Hq/Hkv query heads. Same inputs and output as qspec_decode_attention (qkv holds the rotated
 * queries, the cache already contains this pass's keys); B = number of sequences the cache holds (seq_ids < B).
 * fp32 math; agrees with qspec_decode_attention to fp32 rounding, not bit for bit. qspec_seq_attention_supported():
 * D == 128, Hq/Hkv in {1,2,4,5,8}, T <= 1024, S small enough for shared memory (~1500 positions at Hq/Hkv = 4). */
int qspec_seq_attention_supported(int T, int Hq, int Hkv, int D, int S);
int qspec_seq_attention(const void* qkv, const void* cache_k, const void* cache_v, const int* positions,
                        const int* seq_ids, void* out, int T, int B, int Hq, int Hkv, int D, int S, float sm_scale,
                        void* stream);

/* The attention block between qkv_proj and o_proj as one launch for decode-sized passes: rotary embedding and
 * cache append (qspec_rope_kvcache), attention over the cache (qspec_decode_attention), the o_proj rotation over
 * heads (qspec_hadamard_heads) and, when out_q is given, the per-token quantiser -- quarot_llama.py:207-238.
 * qkv is NOT modified (the separate kernels rotate q in place). attn_out [T, Hq, D] f16 is optional. Outputs are
 * bit-identical to the chain of the separate calls. qspec_attn_block_supported() returns 1 for the shapes it takes
 * (D == 128, Hq/Hkv in {2,4,8}, Hkv in {1,2,4,8}, T and S small enough for shared memory); otherwise use the chain.
 * Tokens of one sequence within a pass must have distinct positions. */
int qspec_attn_block_supported(int T, int Hq, int Hkv, int D, int S);
int qspec_attn_block(const void* qkv, const void* cos_sin_cache, const int* positions, const int* seq_ids,
                     void* cache_k, void* cache_v, void* attn_out, void* out_f16, int8_t* out_q, void* scale_out,
                     int T, int Hq, int Hkv, int D, int S, float sm_scale, float had_scale, float clip_ratio,
                     void* stream);

/* Greedy sampling over the lm_head logits [rows, V] f16 (row stride in elements): out_idx[r] = lowest index of the
 * row maximum (int64, like torch.argmax), out_val[r] (optional) the maximum as f32 -- what LlamaForCausalLM's greedy
 * sampler returns for QSpec's draft/verify token comparison (worker.py acceptance loop). workspace: zero-initialised
 * once by the caller (qspec_argmax_rows_workspace_bytes(rows) bytes), left clean by every call. */
size_t qspec_argmax_rows_workspace_bytes(int rows);
int qspec_argmax_rows(const void* logits, int rows, int V, long long row_stride, long long* out_idx, float* out_val,
                      void* workspace, size_t workspace_bytes, void* stream);

/* Allocation-free variant of qspec_argmax_rows. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, outputs and result as qspec_argmax_rows, but the scratch (per-row arrival counters + slice results) is a
 * file-scope device array inside the library, sized for qspec_argmax_rows_static_max_rows() = 1024 rows (520 KB, held by
 * every GPU that loads the library until the process exits); more rows -> QSPEC_ERR_UNSUPPORTED with a message naming the
 * limit. The call performs no cudaMalloc / cudaFree and calls no allocating library routine: one cudaMemsetAsync of the
 * counters it is about to use and one kernel launch on `stream`. All calls on a GPU share the array (which keeps the last
 * call's contents): calls in stream order are safe, calls that OVERLAP on one GPU (different streams) race. */
int qspec_argmax_rows_static_max_rows(void);
int qspec_argmax_rows_static(const void* logits, int rows, int V, long long row_stride, long long* out_idx, float* out_val,
                             void* stream);

/* ---- Qwen2.5-14B variant of the path (vllm/model_executor/models/qwen2_quarot.py) -------------
 * hidden 5120 = 40 heads x 128, intermediate 13824 = 108 x 128, bias in qkv_proj (the `bias` argument of both GEMMs).
 *
 * down_proj rotation had108 (x) H_128, optionally fused with SiLU*mul in front and the Quantizer behind. Replaces
 * hadamard_utils.matmul_hadU_cuda(x, had108, 108) as called by Qwen2MLP.forward (qwen2_quarot.py:119-127;
 * third-party/QuaRot/quarot/functional/hadamard.py:116-126: FHT_128 * scale in fp32, f16 round, had108 @ [rows,108,128]
 * with fp32 accumulation in ascending k, f16 out) and vLLM's SiluAndMul.
 *   x [rows, 13824] f16 when gate_up == NULL; gate_up [rows, 2*13824] f16, gate_first != 0: [gate ; up] (vLLM's
 *   merged gate_up_proj order, qwen2_quarot.py:80-86), else [up ; gate]. n must be 128. */
int qspec_hadamard_k108(const void* x, const void* gate_up, int gate_first, void* out_f16, int8_t* out_q,
                        void* scale_out, int rows, int n, float scale, float clip_ratio, void* stream);

/* o_proj rotation across 40 heads: out[h, d] = f16(f16(sum_k had40[h][k] x[k, d]) * scale), optional Quantizer.
 * Replaces Qwen2Attention.o_had_trans for o_K = 40 (qwen2_quarot.py:269-277: f16 matmul, then the multiplication by
 * o_had_scale = 1/sqrt(40) rounds a second time) and the Quantizer call (:261). x, out_f16 [rows, 40, 128]. */
int qspec_hadamard_heads40(const void* x, void* out_f16, int8_t* out_q, void* scale_out, int rows, int H, int D,
                           float scale, float clip_ratio, void* stream);

/* qspec_rope_kvcache plus the per-head Hadamard of q and k that Qwen2Attention.qk_had_trans applies after the rotary
 * embedding (qwen2_quarot.py:279-288: fp32 FHT_128 of the f16 rope output, * had_scale = 1/sqrt(128), f16): q in
 * place, k in place and into the cache, v into the cache. D == 128. */
int qspec_rope_qkhad_kvcache(void* qkv, const void* cos_sin_cache, const int* positions, const int* seq_ids,
                             void* cache_k, void* cache_v, int T, int Hq, int Hkv, int D, int S, float had_scale,
                             void* stream);

#ifdef __cplusplus
}
#endif
#endif /* QSPEC_B200_H_ */
