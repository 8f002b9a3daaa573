// lp_common.cuh -- device helpers shared by the sm_100a kernels of liblp_b200.so.
//
// Arithmetic contract (week-2/diet/simplex.py):
//   * every product and every sum is rounded separately (simplex.py:188 `opt(a) + factor*opt(c)`
//     is two Python float operations) -> __dmul_rn / __dadd_rn, never an FMA;
//   * every produced value is snapped: |v| <= tol -> +0.0 (simplex.py:29-34, 191-193, 336, 344, 415);
//   * the ratio test divides (simplex.py:414), the pivot row is scaled by -1*(1.0/pivot)
//     (simplex.py:328) -> IEEE division via __ddiv_rn.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace lpb {

constexpr int kSetUnused = -1;
constexpr int kSetDummy = -2;
constexpr int kNoneInitial = -1;  // basic row never rewritten: every slack column is None
constexpr unsigned long long kKeyNone = ~0ull;

__device__ __forceinline__ double snap(double v, double tol) { return fabs(v) <= tol ? 0.0 : v; }

// a + f*r with two roundings, then snapped (Constraint.sexp, simplex.py:185-194)
__device__ __forceinline__ double axpy_snap(double a, double f, double r, double tol)
{
    return snap(__dadd_rn(a, __dmul_rn(f, r)), tol);
}

// -1 * (1.0 / pivot), snapped (make_pivot_constraint, simplex.py:328-330)
__device__ __forceinline__ double pivot_inverse(double piv, double tol)
{
    return snap(-__ddiv_rn(1.0, piv), tol);
}

// Typed constraint matrices (include/lp_b200.h LPB200_DTYPE_*): element k, or elements 4q..4q+3
// with one streaming load (base + 4q elements must be aligned to four elements), widened to
// fp64 -- exact for every supported type.  Dtype codes are spelled out to keep this header free
// of the public one: F32 = 1, I32 = 2, I16 = 3, I8 = 4.
__device__ __forceinline__ double load_typed1(const char *base, int dtype, size_t k)
{
    switch (dtype) {
    case 1: return (double)__ldcs(reinterpret_cast<const float *>(base) + k);
    case 2: return (double)__ldcs(reinterpret_cast<const int *>(base) + k);
    case 3: return (double)__ldcs(reinterpret_cast<const short *>(base) + k);
    default: return (double)__ldcs(reinterpret_cast<const signed char *>(base) + k);
    }
}

__device__ __forceinline__ void load_typed4(const char *base, int dtype, size_t q, double v[4])
{
    if (dtype == 1) {
        const float4 t = __ldcs(reinterpret_cast<const float4 *>(base) + q);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else if (dtype == 2) {
        const int4 t = __ldcs(reinterpret_cast<const int4 *>(base) + q);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else if (dtype == 3) {
        const short4 t = __ldcs(reinterpret_cast<const short4 *>(base) + q);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
        const char4 t = __ldcs(reinterpret_cast<const char4 *>(base) + q);
        v[0] = (signed char)t.x; v[1] = (signed char)t.y; v[2] = (signed char)t.z; v[3] = (signed char)t.w;
    }
}

// order-preserving map double -> uint64 (no NaNs expected; -0.0 never reaches here: snapped)
__device__ __forceinline__ unsigned long long order_key(double v)
{
    unsigned long long u = (unsigned long long)__double_as_longlong(v);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}

// lexicographic (key, id) minimum across a warp
__device__ __forceinline__ void warp_argmin(unsigned long long &key, int &id)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const unsigned long long ok = __shfl_xor_sync(0xffffffffu, key, off);
        const int oi = __shfl_xor_sync(0xffffffffu, id, off);
        if (ok < key || (ok == key && oi < id)) {
            key = ok;
            id = oi;
        }
    }
}

// ---------------------------------------------------------------------------------------
// CPython 3.12 `set` of small ints (hash(k) == k), as used by the reference for
// `independent` / `dependent` (simplex.py:289-292, 369-381).  The iteration order of
// `independent` IS the entering rule (simplex.py:390-392), so the open-addressing table is
// emulated exactly: LINEAR_PROBES = 9, PERTURB_SHIFT = 5, last-dummy-wins reuse, rebuild at
// fill*5 >= mask*3 to the first power of two > 4*used, re-inserted in old slot order.
// Tables hold int16 keys (variable ids < 32000), kSetUnused or kSetDummy.
// One thread drives these; they are off the critical path (overlapped with the sweep).
// ---------------------------------------------------------------------------------------
struct PySet {
    int size;  // power of two
    int fill;  // active + dummy
    int used;  // active
};

template <typename TableT>
__device__ __forceinline__ void pyset_init_range(TableT *tab, PySet &s, int size, int lo, int hi)
{
    // set(range(lo, hi)): after construction every key sits in its home slot key & mask
    // (contiguous keys never collide while used <= size), fill == used == hi - lo.
    for (int i = 0; i < size; ++i) tab[i] = (TableT)kSetUnused;
    for (int k = lo; k < hi; ++k) tab[k & (size - 1)] = (TableT)k;
    s.size = size;
    s.fill = hi - lo;
    s.used = hi - lo;
}

// Returns the slot the key was stored in; *need_rebuild is set when CPython would resize.
template <typename TableT>
__device__ __forceinline__ int pyset_add(TableT *tab, PySet &s, int key, bool *need_rebuild)
{
    const unsigned mask = (unsigned)s.size - 1u;
    unsigned i = (unsigned)key & mask;
    unsigned perturb = (unsigned)key;
    int freeslot = -1;
    *need_rebuild = false;
    for (;;) {
        const int probes = (i + 9u <= mask) ? 9 : 0;
        for (int j = 0; j <= probes; ++j) {
            const int e = (int)tab[i + j];
            if (e == kSetUnused) {
                if (freeslot >= 0) {
                    tab[freeslot] = (TableT)key;
                    s.used++;
                    return freeslot;
                }
                tab[i + j] = (TableT)key;
                s.fill++;
                s.used++;
                *need_rebuild = ((unsigned)s.fill * 5u >= mask * 3u);
                return (int)(i + j);
            }
            if (e == kSetDummy) freeslot = (int)(i + j);
        }
        perturb >>= 5;
        i = (i * 5u + 1u + perturb) & mask;
    }
}

template <typename TableT>
__device__ __forceinline__ void pyset_remove_slot(TableT *tab, PySet &s, int slot)
{
    tab[slot] = (TableT)kSetDummy;  // fill unchanged
    s.used--;
}

template <typename TableT>
__device__ __forceinline__ void pyset_insert_clean(TableT *tab, unsigned mask, int key)
{
    unsigned perturb = (unsigned)key;
    unsigned i = (unsigned)key & mask;
    for (;;) {
        if ((int)tab[i] == kSetUnused) { tab[i] = (TableT)key; return; }
        if (i + 9u <= mask) {
            for (unsigned j = 1; j <= 9u; ++j)
                if ((int)tab[i + j] == kSetUnused) { tab[i + j] = (TableT)key; return; }
        }
        perturb >>= 5;
        i = (i * 5u + 1u + perturb) & mask;
    }
}

// set_table_resize(so, used > 50000 ? used*2 : used*4): rebuild into `scratch`, copy back.
// Single-thread version (used by the one-CTA-per-LP kernel's bookkeeping lane).
template <typename TableT>
__device__ __forceinline__ void pyset_rebuild(TableT *tab, TableT *scratch, PySet &s)
{
    const int minused = s.used > 50000 ? s.used * 2 : s.used * 4;
    int newsize = 8;
    while (newsize <= minused) newsize <<= 1;
    for (int i = 0; i < newsize; ++i) scratch[i] = (TableT)kSetUnused;
    for (int i = 0; i < s.size; ++i) {
        const int k = (int)tab[i];
        if (k >= 0) pyset_insert_clean(scratch, (unsigned)newsize - 1u, k);
    }
    for (int i = 0; i < newsize; ++i) tab[i] = scratch[i];
    s.size = newsize;
    s.fill = s.used;
}

}  // namespace lpb
