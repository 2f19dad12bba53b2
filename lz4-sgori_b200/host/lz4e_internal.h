/*
 * lz4e_internal.h — calls between the layers INSIDE liblz4e_b200.so that are
 * not part of the public C ABI (include/lz4e_b200.h).
 */
#ifndef LZ4E_INTERNAL_H
#define LZ4E_INTERNAL_H

#include <stddef.h>

#include "../../include/lz4e_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

enum { LZ4E_ATTACH_HOST_WS = 1, LZ4E_ATTACH_CHUNK_POOL = 2 };

/*
 * Per-context state of a layer above the ABI.  The first call for `slot`
 * allocates `bytes` zeroed bytes and remembers `dtor`; later calls return the
 * same pointer (bytes == 0: only look up).  lz4e_b200_ctx_destroy() calls
 * dtor(ctx, p) for every attachment, newest first, BEFORE it releases the
 * context's own streams and arenas, then frees p.
 */
void *lz4e_ctx_attach(lz4e_b200_ctx *ctx, int slot, size_t bytes,
		      void (*dtor)(lz4e_b200_ctx *, void *));

#ifdef __cplusplus
}
#endif
#endif
