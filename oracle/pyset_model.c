/*
 * oracle/pyset_model.c -- TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into
 * the product library).
 *
 * Model of CPython 3.12's `set` for small non-negative int keys (hash(k) == k),
 * restated from the published algorithm of CPython 3.12 Objects/setobject.c
 * (set_add_entry / set_insert_clean / set_table_resize / set_discard_entry /
 * set_next).  CPython is the interpreter the reference runs on; it is NOT part of
 * /root/reference, so parity is anchored on the reference's call sites:
 *
 *   self.independent = set(range(n, nvars))          week-2/diet/simplex.py:289
 *   self.dependent   = set(range(0, n))              week-2/diet/simplex.py:292
 *   s.remove(rem); s.add(add)                        week-2/diet/simplex.py:369-381
 *   for (i, v) in zip_set(in_set): first positive    week-2/diet/simplex.py:140-149, 390-392
 *   for idx in aux_sf.dependent                      week-2/diet/simplex.py:538
 *   aux_sf.independent.remove(nvars)                 week-2/diet/simplex.py:574
 *
 * and is pinned in tests/test_pyset_model.py against the real builtin `set` of the
 * interpreter running the tests (CPython 3.12.3 on the build image and the GPU box).
 */
#include <stdlib.h>
#include <string.h>
#include "lp_oracle.h"

#define PYSET_MINSIZE 8
#define LINEAR_PROBES 9
#define PERTURB_SHIFT 5

void pyset_init(pyset_t *s)
{
    s->size = PYSET_MINSIZE;
    s->table = (int32_t *)malloc(sizeof(int32_t) * (size_t)s->size);
    for (int i = 0; i < s->size; ++i) s->table[i] = PYSET_UNUSED;
    s->fill = 0;
    s->used = 0;
}

void pyset_free(pyset_t *s)
{
    free(s->table);
    s->table = NULL;
    s->size = 0;
}

/* set_insert_clean: table has no dummies and does not contain key. */
static void insert_clean(int32_t *table, uint64_t mask, int32_t key)
{
    uint64_t perturb = (uint64_t)key;
    uint64_t i = (uint64_t)key & mask;
    for (;;) {
        if (table[i] == PYSET_UNUSED) { table[i] = key; return; }
        if (i + LINEAR_PROBES <= mask) {
            for (uint64_t j = 1; j <= LINEAR_PROBES; ++j) {
                if (table[i + j] == PYSET_UNUSED) { table[i + j] = key; return; }
            }
        }
        perturb >>= PERTURB_SHIFT;
        i = (i * 5 + 1 + perturb) & mask;
    }
}

/* set_table_resize(so, minused): rebuild in OLD SLOT ORDER, dropping dummies. */
static void resize(pyset_t *s, int64_t minused)
{
    int64_t newsize = PYSET_MINSIZE;
    while (newsize <= minused) newsize <<= 1;
    int32_t *nt = (int32_t *)malloc(sizeof(int32_t) * (size_t)newsize);
    for (int64_t i = 0; i < newsize; ++i) nt[i] = PYSET_UNUSED;
    for (int i = 0; i < s->size; ++i)
        if (s->table[i] >= 0) insert_clean(nt, (uint64_t)newsize - 1, s->table[i]);
    free(s->table);
    s->table = nt;
    s->size = (int)newsize;
    s->fill = s->used;
}

/* set_add_entry.  Returns 1 if the table was rebuilt (every slot index changed). */
int pyset_add(pyset_t *s, int32_t key)
{
    uint64_t mask = (uint64_t)s->size - 1;
    uint64_t i = (uint64_t)key & mask;
    uint64_t perturb = (uint64_t)key;
    int64_t freeslot = -1;
    for (;;) {
        int probes = (i + LINEAR_PROBES <= mask) ? LINEAR_PROBES : 0;
        for (int j = 0; j <= probes; ++j) {
            int32_t e = s->table[i + (uint64_t)j];
            if (e == PYSET_UNUSED) {
                if (freeslot >= 0) {          /* reuse the LAST dummy seen on the way */
                    s->table[freeslot] = key;
                    s->used++;
                    return 0;
                }
                s->table[i + (uint64_t)j] = key;
                s->fill++;
                s->used++;
                if ((uint64_t)s->fill * 5 < mask * 3) return 0;
                resize(s, s->used > 50000 ? (int64_t)s->used * 2 : (int64_t)s->used * 4);
                return 1;
            }
            if (e == key) return 0;           /* already present */
            if (e == PYSET_DUMMY) freeslot = (int64_t)(i + (uint64_t)j);
        }
        perturb >>= PERTURB_SHIFT;
        i = (i * 5 + 1 + perturb) & mask;
    }
}

/* set_discard_entry via set_lookkey: the slot becomes a dummy; fill is unchanged. */
int pyset_remove(pyset_t *s, int32_t key)
{
    uint64_t mask = (uint64_t)s->size - 1;
    uint64_t i = (uint64_t)key & mask;
    uint64_t perturb = (uint64_t)key;
    for (;;) {
        int probes = (i + LINEAR_PROBES <= mask) ? LINEAR_PROBES : 0;
        for (int j = 0; j <= probes; ++j) {
            int32_t e = s->table[i + (uint64_t)j];
            if (e == PYSET_UNUSED) return -1;  /* KeyError */
            if (e == key) {
                s->table[i + (uint64_t)j] = PYSET_DUMMY;
                s->used--;
                return 0;
            }
        }
        perturb >>= PERTURB_SHIFT;
        i = (i * 5 + 1 + perturb) & mask;
    }
}

/* set(range(lo, hi)): empty 8-slot table, then add one key at a time. */
void pyset_init_range(pyset_t *s, int32_t lo, int32_t hi)
{
    pyset_init(s);
    for (int32_t k = lo; k < hi; ++k) pyset_add(s, k);
}

/* Iteration order = ascending slot index over active entries (set_next). */
int pyset_order(const pyset_t *s, int32_t *out)
{
    int n = 0;
    for (int i = 0; i < s->size; ++i)
        if (s->table[i] >= 0) out[n++] = s->table[i];
    return n;
}

/* ---- flat test entry point: replay a remove/add script, report order + size ---- */
int lp_oracle_pyset_replay(int32_t lo, int32_t hi, int nops, const int32_t *ops /* [nops][2] = (remove, add) */,
                           int32_t *order_out, int32_t *table_size_out)
{
    pyset_t s;
    pyset_init_range(&s, lo, hi);
    for (int t = 0; t < nops; ++t) {
        if (ops[2 * t] >= 0 && pyset_remove(&s, ops[2 * t]) != 0) { pyset_free(&s); return -1; }
        if (ops[2 * t + 1] >= 0) pyset_add(&s, ops[2 * t + 1]);
    }
    int cnt = pyset_order(&s, order_out);
    *table_size_out = s.size;
    pyset_free(&s);
    return cnt;
}
