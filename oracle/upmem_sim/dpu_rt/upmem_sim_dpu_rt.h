/*
 * upmem_sim_dpu_rt.h -- DPU-side runtime stand-in (TEST INFRASTRUCTURE, see oracle/README.md).
 *
 * The closed UPMEM SDK is not in this image.  To run the reference's DPU program
 * (/root/reference/upmem/src/dpu/emb_dpu_lookup.c) *verbatim* on the host CPU, the five
 * DPU runtime headers it includes (mram.h, perfcounter.h, alloc.h, defs.h, sem.h) resolve
 * to this file.  It models only what that program can observe:
 *   - MRAM and WRAM variables are ordinary globals (8-byte aligned, as on the DPU);
 *   - mram_read/mram_write are memcpy with the MRAM address aligned DOWN to 8 bytes
 *     (the DMA engine ignores the low three address bits -- this is why the reference reads
 *     8 bytes at &emb_data[ind] and then picks lane ind % 2, emb_dpu_lookup.c:112-114);
 *   - tasklets are run one after another by the host-side simulator, so semaphores and
 *     the perf counter are no-ops and me() returns the id the simulator set.
 * Written from the public UPMEM programming-model documentation; no SDK code.
 */
#ifndef UPMEM_SIM_DPU_RT_H
#define UPMEM_SIM_DPU_RT_H

#include <stdint.h>
#include <string.h>

#define __mram_noinit __attribute__((aligned(8)))
#define __mram __attribute__((aligned(8)))
#define __mram_ptr
#define __host __attribute__((aligned(8)))
#define __dma_aligned __attribute__((aligned(8)))

/* set by the simulator before each tasklet run (see host_sdk/upmem_sim_host.c) */
extern unsigned int upmem_sim_current_tasklet;

static inline unsigned int me(void) { return upmem_sim_current_tasklet; }

static inline void mram_read(const void *from_mram, void *to_wram, unsigned int nb_bytes)
{
    const char *src = (const char *)((uintptr_t)from_mram & ~(uintptr_t)7);
    memcpy(to_wram, src, nb_bytes);
}

static inline void mram_write(const void *from_wram, void *to_mram, unsigned int nb_bytes)
{
    char *dst = (char *)((uintptr_t)to_mram & ~(uintptr_t)7);
    memcpy(dst, from_wram, nb_bytes);
}

/* alloc.h */
static inline void mem_reset(void) {}

/* perfcounter.h */
typedef enum { COUNT_CYCLES = 0, COUNT_INSTRUCTIONS = 1 } perfcounter_config_t;
static inline void perfcounter_config(perfcounter_config_t cfg, int reset) { (void)cfg; (void)reset; }
static inline uint32_t perfcounter_get(void) { return 0; }

/* sem.h: sequential tasklets never contend */
typedef int upmem_sim_sem_t;
#define SEMAPHORE_INIT(_name, _count) upmem_sim_sem_t _name = (_count)
static inline void sem_take(upmem_sim_sem_t *s) { (void)s; }
static inline void sem_give(upmem_sim_sem_t *s) { (void)s; }

#endif
