// Short-dialogue attention fast path, 2 bf16 plane(s): see attention_mma_impl.cuh
#define M2F_ATT_P 2
#include "attention_mma_impl.cuh"
