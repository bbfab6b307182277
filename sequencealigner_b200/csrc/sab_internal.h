// sab_internal.h -- shared between the translation units of libseqalign_b200.so
#pragma once
#include "../../include/seqalign_b200.h"

// device ordinal of the context, or -1 when it has no usable device
int sab_internal_device(const sab_ctx *ctx);
// records the message in the context and returns `code`
int sab_internal_fail(sab_ctx *ctx, int code, const char *fmt, ...);
