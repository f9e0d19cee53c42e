/*
 * dpu.h -- host-side stand-in for the UPMEM SDK's <dpu.h>.  TEST INFRASTRUCTURE
 * (oracle/README.md): lets the reference's host library (/root/reference/upmem/include/
 * emb_host.h, upmem/src/load_generator.c) compile and run verbatim on a CPU-only box.
 *
 * Only the entry points those two files use are modelled, with the documented SDK
 * behaviour that they can observe:
 *   dpu_alloc        N DPUs, grouped into ranks of 64 in allocation order
 *   dpu_load         loads a "DPU binary": here a shared object built from the verbatim
 *                    DPU source by dpu_rt/dpu_program_unit.c
 *   DPU_FOREACH      iterates DPUs in allocation order; the optional index variable is
 *                    incremented in ITS OWN TYPE (so a uint8_t index wraps at 256, which
 *                    the reference relies on / suffers from, emb_host.h:163-172)
 *   dpu_prepare_xfer / dpu_push_xfer   per-DPU buffers, DPUs without a prepared buffer
 *                    are skipped; transfers address MRAM/host symbols by name
 *   dpu_launch       runs every DPU of the set, tasklets one after another
 *   dpu_callback     calls fn(rank_set, rank_index, arg) once per rank
 *   dpu_sync, dpu_copy_from, DPU_ASSERT
 * Written from the public SDK documentation; contains no SDK code.
 */
#ifndef UPMEM_SIM_DPU_H
#define UPMEM_SIM_DPU_H

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { DPU_OK = 0, DPU_ERR_SIM = 1 } dpu_error_t;
typedef enum { DPU_XFER_TO_DPU = 0, DPU_XFER_FROM_DPU = 1 } dpu_xfer_t;
typedef enum { DPU_XFER_DEFAULT = 0 } dpu_xfer_flags_t;
typedef enum { DPU_SYNCHRONOUS = 0, DPU_ASYNCHRONOUS = 1 } dpu_launch_policy_t;
typedef enum { DPU_CALLBACK_DEFAULT = 0, DPU_CALLBACK_ASYNC = 1 } dpu_callback_flags_t;

#define DPU_ALLOCATE_ALL 0xffffffffu
#define UPMEM_SIM_SET_MAGIC 0x55504d53u /* 'UPMS' */

struct upmem_sim_system;

/* A set is a contiguous [first, first+count) window of one allocation. */
struct dpu_set_t {
    uint32_t magic;
    uint32_t first;
    uint32_t count;
    struct upmem_sim_system *sys;
};

struct upmem_sim_iter {
    struct dpu_set_t set;
    uint32_t pos;
    int once;
};

dpu_error_t dpu_alloc(uint32_t nr_dpus, const char *profile, struct dpu_set_t *set);
dpu_error_t dpu_free(struct dpu_set_t set);
dpu_error_t dpu_load(struct dpu_set_t set, const char *binary_path, void **program);
dpu_error_t dpu_prepare_xfer(struct dpu_set_t set, void *buffer);
dpu_error_t dpu_push_xfer(struct dpu_set_t set, dpu_xfer_t xfer, const char *symbol_name,
                          uint32_t symbol_offset, size_t length, dpu_xfer_flags_t flags);
dpu_error_t dpu_copy_from(struct dpu_set_t set, const char *symbol_name, uint32_t symbol_offset,
                          void *dst, size_t length);
dpu_error_t dpu_copy_to(struct dpu_set_t set, const char *symbol_name, uint32_t symbol_offset,
                        const void *src, size_t length);
dpu_error_t dpu_launch(struct dpu_set_t set, dpu_launch_policy_t policy);
dpu_error_t dpu_sync(struct dpu_set_t set);
dpu_error_t dpu_callback(struct dpu_set_t set,
                         dpu_error_t (*fn)(struct dpu_set_t, uint32_t, void *), void *arg,
                         dpu_callback_flags_t flags);
dpu_error_t dpu_get_nr_dpus(struct dpu_set_t set, uint32_t *nr_dpus);
dpu_error_t dpu_get_nr_ranks(struct dpu_set_t set, uint32_t *nr_ranks);
const char *dpu_error_to_string(dpu_error_t e);

/* simulator controls (not part of the SDK surface) */
void upmem_sim_set_tasklet_order(const unsigned int *order, unsigned int n);
unsigned long upmem_sim_launch_count(void);

struct upmem_sim_iter upmem_sim_iter_make(const struct dpu_set_t *set);
int upmem_sim_iter_load(struct upmem_sim_iter *it, struct dpu_set_t *one);

#define UPMEM_SIM_FOREACH_N(_set, _one)                                                       \
    for (struct upmem_sim_iter _upmem_it = upmem_sim_iter_make(&(_set)); _upmem_it.once;      \
         _upmem_it.once = 0)                                                                  \
        for (; upmem_sim_iter_load(&_upmem_it, &(_one)); _upmem_it.pos++)

#define UPMEM_SIM_FOREACH_I(_set, _one, _i)                                                   \
    for (struct upmem_sim_iter _upmem_it = upmem_sim_iter_make(&(_set)); _upmem_it.once;      \
         _upmem_it.once = 0)                                                                  \
        for ((_i) = 0; upmem_sim_iter_load(&_upmem_it, &(_one)); _upmem_it.pos++, (_i)++)

#define UPMEM_SIM_PICK(_1, _2, _3, _name, ...) _name
#define DPU_FOREACH(...)                                                                      \
    UPMEM_SIM_PICK(__VA_ARGS__, UPMEM_SIM_FOREACH_I, UPMEM_SIM_FOREACH_N, 0)(__VA_ARGS__)

#define DPU_ASSERT(_stmt)                                                                     \
    do {                                                                                      \
        dpu_error_t _upmem_e = (_stmt);                                                       \
        if (_upmem_e != DPU_OK) {                                                             \
            fprintf(stderr, "%s:%d(%s): DPU Error (%s)\n", __FILE__, __LINE__, __func__,      \
                    dpu_error_to_string(_upmem_e));                                           \
            exit(EXIT_FAILURE);                                                               \
        }                                                                                     \
    } while (0)

#ifdef __cplusplus
}
#endif
#endif
