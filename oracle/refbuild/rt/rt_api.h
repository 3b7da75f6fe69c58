/* ORACLE build support (test infrastructure, not product code).
 *
 * Minimal host stand-in for the PULP-SDK runtime header "rt/rt_api.h" so that the UNMODIFIED
 * reference sources under /root/reference/edge-eegnet_wolf/src/cl/{net,func}/ compile with gcc on
 * x86-64 (recipe: SURVEY.md Appendix B.3).  Nothing here is copied from the reference; it provides
 * only the symbols those sources use: the XpulpV2 builtins (__SUMDOTP4, __PACK4, __AND4, __CLIP_R,
 * __MAX, __MAC), the L1/L2 allocator, the cluster DMA (a memcpy), and the 8-core team fork (a serial
 * loop).  Pointers are passed through `unsigned int` by the reference (e.g. layer1.c:137-138), so the
 * harness is linked -no-pie and every buffer comes from a MAP_32BIT arena (shim.c).
 */
#ifndef ORACLE_RT_API_SHIM_H
#define ORACLE_RT_API_SHIM_H

#include <stdint.h>
#include <stddef.h>
#include <stdio.h>
#include <string.h>

typedef signed char v4s __attribute__((vector_size(4)));

#define RT_L2_DATA
#define RT_CL_DATA
#define RT_LOCAL_DATA
#define RT_FC_SHARED_DATA
#define RT_FC_GLOBAL_DATA

#define RT_ALLOC_CL_DATA 0
#define RT_ALLOC_L2_CL_DATA 1
#define RT_ALLOC_FC_DATA 2
#define RT_DMA_DIR_EXT2LOC 0
#define RT_DMA_DIR_LOC2EXT 1

typedef int rt_dma_copy_t;

void* rt_alloc(int where, int size);
void rt_free(int where, void* p, int size);

static inline void rt_dma_memcpy(unsigned int ext, unsigned int loc, unsigned short size, int dir,
                                 int merge, rt_dma_copy_t* copy) {
    (void)merge; (void)copy;
    if (dir == RT_DMA_DIR_EXT2LOC) memcpy((void*)(uintptr_t)loc, (const void*)(uintptr_t)ext, size);
    else memcpy((void*)(uintptr_t)ext, (const void*)(uintptr_t)loc, size);
}
static inline void rt_dma_wait(rt_dma_copy_t* copy) { (void)copy; }

extern int oracle_shim_core_id;
static inline int rt_core_id(void) { return oracle_shim_core_id; }
static inline void rt_team_barrier(void) {}
static inline void rt_team_fork(int n, void (*fn)(void*), void* arg) {
    for (int i = 0; i < n; i++) { oracle_shim_core_id = i; fn(arg); }
    oracle_shim_core_id = 0;
}

/* builtins: functions, not macros -- call sites pass expressions with side effects (layer2.c:106) */
static inline int32_t __CLIP_R(int32_t x, int32_t b) { return x < -(b + 1) ? -(b + 1) : (x > b ? b : x); }
static inline int32_t __MAX(int32_t a, int32_t b) { return a > b ? a : b; }
static inline int32_t __MAC(int32_t acc, int32_t a, int32_t b) { return acc + a * b; }
static inline int32_t __SUMDOTP4(v4s a, v4s b, int32_t acc) {
    return acc + a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3];
}
static inline v4s __PACK4(int a, int b, int c, int d) {
    return (v4s){(signed char)a, (signed char)b, (signed char)c, (signed char)d};
}
static inline v4s __AND4(v4s a, v4s b) { return a & b; }

#endif
