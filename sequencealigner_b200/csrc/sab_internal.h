// sab_internal.h -- shared between the translation units of libseqalign_b200.so
#pragma once
#include "../../include/seqalign_b200.h"

#include <vector>

// sab_config.flags bit set by sab_multi_* on its contexts (not part of the public flag set): the
// devices pull row strips from a shared cursor, so the panels keep a uniform height (panel_bounds)
enum { SAB_FLAG_INTERNAL_DYNAMIC = 1 << 30 };

// device ordinal of the context, or -1 when it has no usable device
int sab_internal_device(const sab_ctx *ctx);
// records the message in the context and returns `code`
int sab_internal_fail(sab_ctx *ctx, int code, const char *fmt, ...);
// per-context scratch of sab_filter.cu (device buffers kept between calls); `release` is what
// sab_destroy calls on it
void **sab_internal_filter_slot(sab_ctx *ctx, void (*release)(void *));
// sab_config.flags of the context
void sab_internal_flags(const sab_ctx *ctx, int *flags);
// where sab_filter* report their wall time / match count (sab_stats.filter_ms, .filter_matches)
void sab_internal_filter_stats(sab_ctx *ctx, double ms, int64_t matches);

// Phase A of the filter on this context's device: every pair i<j of shard `shard` of `nshards`
// whose prefix identity reaches the threshold, as (j << 32 | i), in no particular order.
// *overflow is set (and `pairs` left incomplete) when the device-side list ran full.
int sab_internal_filter_matches(sab_ctx *ctx, const uint8_t *seqs, const sab_meta *meta, int32_t num,
				float threshold, int shard, int nshards, std::vector<uint64_t> &pairs,
				int *overflow);
