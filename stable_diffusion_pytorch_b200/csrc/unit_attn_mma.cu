// Translation unit 2: the mma.sync attention family (dense d > 20, short context, their backward).
#include "attention.cu"
#include "attention_ctx.cu"
#include "attention_bwd.cu"
#include "attention_ctx_bwd.cu"
