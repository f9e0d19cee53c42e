/*
 * upmem_sim_host.c -- functional stand-in for the UPMEM host SDK (see dpu.h in this dir).
 * TEST INFRASTRUCTURE (oracle/README.md); never linked into the product library.
 *
 * Model: one loaded DPU program image (a shared object with ordinary globals for the
 * MRAM / WRAM variables) is time-shared by all simulated DPUs.  Every DPU keeps its own
 * copy of each *exported* symbol (what the host can address: MRAM and __host variables).
 * dpu_launch() swaps a DPU's symbols into the image, runs the tasklets one after another
 * in the configured order, and swaps the symbols back out.  WRAM scratch that the host
 * cannot address is not preserved per DPU; the reference program rewrites all of it on
 * every launch (emb_dpu_lookup.c:42-57) so this is unobservable.
 */
#define _GNU_SOURCE
#include "dpu.h"

#include <dlfcn.h>
#include <string.h>

#define SIM_DPUS_PER_RANK 64u
#define SIM_MAX_SYMS 16
#define SIM_SMALL_SYMBOL (1u << 20) /* symbols up to 1 MiB are snapshotted whole */
#define SIM_MAX_TASKLETS 24

struct sim_program_symbol { /* must match dpu_rt/dpu_program_unit.c */
    const char *name;
    void *addr;
    size_t size;
};

struct sim_store {
    unsigned char *data;
    size_t extent; /* bytes of the symbol this DPU holds, from offset 0 */
};

struct sim_dpu {
    void *pending;
    struct sim_store store[SIM_MAX_SYMS];
};

struct upmem_sim_system {
    uint32_t nr_dpus;
    struct sim_dpu *dpus;
    void *image;
    const struct sim_program_symbol *symtab;
    int nr_syms;
    size_t image_extent[SIM_MAX_SYMS]; /* bytes of each image global last written by a swap-in */
    unsigned char *pristine[SIM_MAX_SYMS]; /* load-time contents of the small symbols (initialisers) */
    int (*entry)(void);
    unsigned int *current_tasklet;
    unsigned int nr_tasklets;
};

static unsigned int g_order[SIM_MAX_TASKLETS];
static unsigned int g_order_n = 0;
static unsigned long g_launches = 0;

void upmem_sim_set_tasklet_order(const unsigned int *order, unsigned int n)
{
    if (n > SIM_MAX_TASKLETS)
        n = SIM_MAX_TASKLETS;
    for (unsigned int i = 0; i < n; i++)
        g_order[i] = order[i];
    g_order_n = n;
}

unsigned long upmem_sim_launch_count(void) { return g_launches; }

const char *dpu_error_to_string(dpu_error_t e) { return e == DPU_OK ? "ok" : "simulator error"; }

static int set_ok(const struct dpu_set_t *s)
{
    return s && s->magic == UPMEM_SIM_SET_MAGIC && s->sys != NULL &&
           (uint64_t)s->first + s->count <= s->sys->nr_dpus;
}

struct upmem_sim_iter upmem_sim_iter_make(const struct dpu_set_t *set)
{
    struct upmem_sim_iter it;
    memset(&it, 0, sizeof it);
    it.once = 1;
    if (set_ok(set))
        it.set = *set; /* an uninitialised set (emb_host.h:250,361) iterates zero times */
    return it;
}

int upmem_sim_iter_load(struct upmem_sim_iter *it, struct dpu_set_t *one)
{
    if (it->set.magic != UPMEM_SIM_SET_MAGIC || it->pos >= it->set.count)
        return 0;
    one->magic = UPMEM_SIM_SET_MAGIC;
    one->sys = it->set.sys;
    one->first = it->set.first + it->pos;
    one->count = 1;
    return 1;
}

dpu_error_t dpu_alloc(uint32_t nr_dpus, const char *profile, struct dpu_set_t *set)
{
    (void)profile;
    if (nr_dpus == 0 || nr_dpus == DPU_ALLOCATE_ALL)
        return DPU_ERR_SIM;
    struct upmem_sim_system *sys = calloc(1, sizeof *sys);
    if (!sys)
        return DPU_ERR_SIM;
    sys->nr_dpus = nr_dpus;
    sys->dpus = calloc(nr_dpus, sizeof *sys->dpus);
    if (!sys->dpus)
        return DPU_ERR_SIM;
    set->magic = UPMEM_SIM_SET_MAGIC;
    set->first = 0;
    set->count = nr_dpus;
    set->sys = sys;
    return DPU_OK;
}

dpu_error_t dpu_free(struct dpu_set_t set)
{
    if (!set_ok(&set))
        return DPU_ERR_SIM;
    struct upmem_sim_system *sys = set.sys;
    for (uint32_t d = 0; d < sys->nr_dpus; d++)
        for (int s = 0; s < SIM_MAX_SYMS; s++)
            free(sys->dpus[d].store[s].data);
    for (int s = 0; s < SIM_MAX_SYMS; s++)
        free(sys->pristine[s]);
    free(sys->dpus);
    free(sys);
    return DPU_OK;
}

dpu_error_t dpu_load(struct dpu_set_t set, const char *binary_path, void **program)
{
    if (!set_ok(&set))
        return DPU_ERR_SIM;
    struct upmem_sim_system *sys = set.sys;
    const char *override = getenv("UPMEM_SIM_DPU_BINARY");
    const char *path = (override && *override) ? override : binary_path;
    sys->image = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!sys->image) {
        fprintf(stderr, "upmem_sim: cannot load DPU program '%s': %s\n", path, dlerror());
        return DPU_ERR_SIM;
    }
    sys->symtab = dlsym(sys->image, "upmem_sim_symbols");
    sys->entry = (int (*)(void))dlsym(sys->image, "upmem_sim_dpu_main");
    sys->current_tasklet = dlsym(sys->image, "upmem_sim_current_tasklet");
    const unsigned int *nt = dlsym(sys->image, "upmem_sim_nr_tasklets");
    if (!sys->symtab || !sys->entry || !sys->current_tasklet || !nt)
        return DPU_ERR_SIM;
    sys->nr_tasklets = *nt;
    sys->nr_syms = 0;
    while (sys->symtab[sys->nr_syms].name)
        sys->nr_syms++;
    if (sys->nr_syms > SIM_MAX_SYMS || sys->nr_tasklets > SIM_MAX_TASKLETS)
        return DPU_ERR_SIM;
    /* every DPU boots from the program's initialised data (e.g. `__host uint8_t first_run = 1`,
     * emb_dpu_lookup.c:34): remember it before any launch dirties the shared image */
    for (int s = 0; s < sys->nr_syms; s++) {
        size_t sz = sys->symtab[s].size;
        if (sz <= SIM_SMALL_SYMBOL) {
            sys->pristine[s] = malloc(sz ? sz : 1);
            if (!sys->pristine[s])
                return DPU_ERR_SIM;
            memcpy(sys->pristine[s], sys->symtab[s].addr, sz);
        }
    }
    if (program)
        *program = sys->image;
    return DPU_OK;
}

static int find_symbol(const struct upmem_sim_system *sys, const char *name)
{
    for (int s = 0; s < sys->nr_syms; s++)
        if (strcmp(sys->symtab[s].name, name) == 0)
            return s;
    return -1;
}

static int store_reserve(struct sim_store *st, size_t extent)
{
    if (extent <= st->extent)
        return 0;
    unsigned char *p = realloc(st->data, extent);
    if (!p)
        return -1;
    memset(p + st->extent, 0, extent - st->extent);
    st->data = p;
    st->extent = extent;
    return 0;
}

dpu_error_t dpu_prepare_xfer(struct dpu_set_t set, void *buffer)
{
    if (!set_ok(&set))
        return DPU_ERR_SIM;
    for (uint32_t d = set.first; d < set.first + set.count; d++)
        set.sys->dpus[d].pending = buffer;
    return DPU_OK;
}

static dpu_error_t xfer_one(struct upmem_sim_system *sys, uint32_t d, int s, dpu_xfer_t xfer,
                            uint32_t offset, void *host, size_t length)
{
    if ((size_t)offset + length > sys->symtab[s].size)
        return DPU_ERR_SIM; /* the SDK rejects out-of-symbol transfers */
    struct sim_store *st = &sys->dpus[d].store[s];
    if (store_reserve(st, (size_t)offset + length) != 0)
        return DPU_ERR_SIM;
    if (xfer == DPU_XFER_TO_DPU)
        memcpy(st->data + offset, host, length);
    else
        memcpy(host, st->data + offset, length);
    return DPU_OK;
}

dpu_error_t dpu_push_xfer(struct dpu_set_t set, dpu_xfer_t xfer, const char *symbol_name,
                          uint32_t symbol_offset, size_t length, dpu_xfer_flags_t flags)
{
    (void)flags;
    if (!set_ok(&set))
        return DPU_ERR_SIM;
    struct upmem_sim_system *sys = set.sys;
    int s = find_symbol(sys, symbol_name);
    if (s < 0)
        return DPU_ERR_SIM;
    dpu_error_t rc = DPU_OK;
    for (uint32_t d = set.first; d < set.first + set.count; d++) {
        struct sim_dpu *dpu = &sys->dpus[d];
        if (dpu->pending) { /* DPUs with no prepared buffer take no part in the transfer */
            if (xfer_one(sys, d, s, xfer, symbol_offset, dpu->pending, length) != DPU_OK)
                rc = DPU_ERR_SIM;
            dpu->pending = NULL;
        }
    }
    return rc;
}

dpu_error_t dpu_copy_from(struct dpu_set_t set, const char *symbol_name, uint32_t symbol_offset,
                          void *dst, size_t length)
{
    if (!set_ok(&set) || set.count != 1)
        return DPU_ERR_SIM;
    int s = find_symbol(set.sys, symbol_name);
    if (s < 0)
        return DPU_ERR_SIM;
    return xfer_one(set.sys, set.first, s, DPU_XFER_FROM_DPU, symbol_offset, dst, length);
}

dpu_error_t dpu_copy_to(struct dpu_set_t set, const char *symbol_name, uint32_t symbol_offset,
                        const void *src, size_t length)
{
    if (!set_ok(&set))
        return DPU_ERR_SIM;
    int s = find_symbol(set.sys, symbol_name);
    if (s < 0)
        return DPU_ERR_SIM;
    for (uint32_t d = set.first; d < set.first + set.count; d++)
        if (xfer_one(set.sys, d, s, DPU_XFER_TO_DPU, symbol_offset, (void *)src, length) != DPU_OK)
            return DPU_ERR_SIM;
    return DPU_OK;
}

static dpu_error_t run_one_dpu(struct upmem_sim_system *sys, uint32_t d)
{
    struct sim_dpu *dpu = &sys->dpus[d];
    /* swap in: this DPU's copy of every host-addressable symbol */
    for (int s = 0; s < sys->nr_syms; s++) {
        struct sim_store *st = &dpu->store[s];
        unsigned char *g = sys->symtab[s].addr;
        size_t full = sys->symtab[s].size;
        if (sys->pristine[s] && st->extent < full) {
            /* first boot of this DPU (or a host write covering only a prefix of the symbol):
             * the rest still holds the program's initial value */
            size_t had = st->extent;
            if (store_reserve(st, full) != 0)
                return DPU_ERR_SIM;
            memcpy(st->data + had, sys->pristine[s] + had, full - had);
        }
        if (st->extent)
            memcpy(g, st->data, st->extent);
        if (sys->image_extent[s] > st->extent) /* clear what another DPU left behind */
            memset(g + st->extent, 0, sys->image_extent[s] - st->extent);
        sys->image_extent[s] = st->extent;
    }
    /* run the tasklets back to back */
    unsigned int order[SIM_MAX_TASKLETS];
    unsigned int n = sys->nr_tasklets;
    for (unsigned int i = 0; i < n; i++)
        order[i] = i;
    if (g_order_n == n)
        memcpy(order, g_order, n * sizeof order[0]);
    for (unsigned int i = 0; i < n; i++) {
        *sys->current_tasklet = order[i];
        (void)sys->entry();
    }
    /* swap out */
    for (int s = 0; s < sys->nr_syms; s++) {
        struct sim_store *st = &dpu->store[s];
        size_t want = sys->symtab[s].size <= SIM_SMALL_SYMBOL ? sys->symtab[s].size : st->extent;
        if (want == 0)
            continue;
        if (store_reserve(st, want) != 0)
            return DPU_ERR_SIM;
        memcpy(st->data, sys->symtab[s].addr, want);
        if (sys->image_extent[s] < want)
            sys->image_extent[s] = want;
    }
    return DPU_OK;
}

dpu_error_t dpu_launch(struct dpu_set_t set, dpu_launch_policy_t policy)
{
    (void)policy; /* asynchronous launches complete before returning as well */
    if (!set_ok(&set) || !set.sys->image)
        return DPU_ERR_SIM;
    g_launches++;
    for (uint32_t d = set.first; d < set.first + set.count; d++)
        if (run_one_dpu(set.sys, d) != DPU_OK)
            return DPU_ERR_SIM;
    return DPU_OK;
}

dpu_error_t dpu_sync(struct dpu_set_t set)
{
    return set_ok(&set) ? DPU_OK : DPU_ERR_SIM;
}

dpu_error_t dpu_callback(struct dpu_set_t set,
                         dpu_error_t (*fn)(struct dpu_set_t, uint32_t, void *), void *arg,
                         dpu_callback_flags_t flags)
{
    (void)flags; /* launches are already complete, so an async callback can run at once */
    if (!set_ok(&set))
        return DPU_ERR_SIM;
    uint32_t first_rank = set.first / SIM_DPUS_PER_RANK;
    uint32_t last_rank = (set.first + set.count - 1) / SIM_DPUS_PER_RANK;
    for (uint32_t r = first_rank; r <= last_rank; r++) {
        struct dpu_set_t rank = set;
        uint32_t lo = r * SIM_DPUS_PER_RANK, hi = lo + SIM_DPUS_PER_RANK;
        if (lo < set.first)
            lo = set.first;
        if (hi > set.first + set.count)
            hi = set.first + set.count;
        rank.first = lo;
        rank.count = hi - lo;
        dpu_error_t e = fn(rank, r - first_rank, arg);
        if (e != DPU_OK)
            return e;
    }
    return DPU_OK;
}

dpu_error_t dpu_get_nr_dpus(struct dpu_set_t set, uint32_t *nr_dpus)
{
    if (!set_ok(&set))
        return DPU_ERR_SIM;
    *nr_dpus = set.count;
    return DPU_OK;
}

dpu_error_t dpu_get_nr_ranks(struct dpu_set_t set, uint32_t *nr_ranks)
{
    if (!set_ok(&set))
        return DPU_ERR_SIM;
    *nr_ranks = (set.first + set.count - 1) / SIM_DPUS_PER_RANK - set.first / SIM_DPUS_PER_RANK + 1;
    return DPU_OK;
}
