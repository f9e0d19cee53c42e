/*
 * dpu_program_unit.c -- wraps one verbatim DPU program source into a loadable image.
 * TEST INFRASTRUCTURE (oracle/README.md).
 *
 * Built as:  gcc -shared -fPIC -Dmain=upmem_sim_dpu_main
 *                -DUPMEM_SIM_DPU_SOURCE='"/root/reference/upmem/src/dpu/emb_dpu_lookup.c"'
 *                -DUPMEM_SIM_SYMBOLS='X(emb_data) X(input_indices) ...'  dpu_program_unit.c
 * The reference source is #included where it lies (never copied); after it, this unit
 * emits the symbol table (name, address, size) the host-side simulator needs to serve
 * dpu_push_xfer()/dpu_copy_from() by symbol name, as the real loader does from the ELF.
 */
#include <stddef.h>

unsigned int upmem_sim_current_tasklet = 0;

#include UPMEM_SIM_DPU_SOURCE

struct upmem_sim_symbol {
    const char *name;
    void *addr;
    size_t size;
};

#define X(_s) { #_s, (void *)&_s, sizeof(_s) },
const struct upmem_sim_symbol upmem_sim_symbols[] = { UPMEM_SIM_SYMBOLS { NULL, NULL, 0 } };
#undef X

const unsigned int upmem_sim_nr_tasklets = NR_TASKLETS;
