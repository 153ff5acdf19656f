// Short-dialogue attention fast path, 3 bf16 plane(s): see attention_mma_impl.cuh
#define M2F_ATT_P 3
#include "attention_mma_impl.cuh"
