/* b200svs - debug / bring-up entry points.  NOT part of the product ABI (include/b200svs.h): nothing on the
 * synthesis path calls these; they are exported by the same library so that hardware facts the kernels rely on stay
 * checkable on a GPU box (tests/test_gpu_misc.py). */
#ifndef B200SVS_DEBUG_H_
#define B200SVS_DEBUG_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* GEMM of a row-shifted shared-memory operand view: pins the UMMA descriptor start-address / base-offset semantics
 * that conv1d_tc's tap loop relies on (one activation tile, every tap a descriptor shifted by tap*dil rows;
 * DESIGN.md 4.1). */
int b200svs_probe_rowshift(const void* a_f16, const void* b_f16, float* out, int32_t rows_a, int32_t shift,
                           int32_t base_offset_mode, void* stream);

/* Diagnostic: per-tile clock64 stamps of b200svs_resblock_pair's pipeline roles into dev_buf[grid][tiles_per_cta][16]
 * (int64); NULL switches it off.  Slot meanings: tools/trace_pair.py. */
int b200svs_resblock_pair_set_trace(void* dev_buf, int32_t tiles_per_cta);

/* Diagnostic: clock64 stamps of the attention_ts roles of CTA (1, 0, 0) into dev_buf[tiles][16] (int64); NULL switches it
 * off.  Slot meanings: tools/trace_attn.py. */
int b200svs_attention_set_trace(void* dev_buf, int32_t tiles);

#ifdef __cplusplus
}
#endif
#endif /* B200SVS_DEBUG_H_ */
