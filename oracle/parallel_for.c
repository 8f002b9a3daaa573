/*
 * oracle/parallel_for.c -- TEST INFRASTRUCTURE ONLY.
 * Minimal pthread work-sharing loop for the timed CPU baseline (the reference itself is
 * single-threaded; the baseline runs one independent problem per host thread).
 */
#include <pthread.h>
#include <stdatomic.h>
#include <stdlib.h>
#include <unistd.h>
#include "lp_oracle.h"

typedef struct {
    atomic_int next;
    int count;
    void (*fn)(void *, int);
    void *arg;
} pool_t;

static void *worker(void *p)
{
    pool_t *pool = (pool_t *)p;
    for (;;) {
        const int i = atomic_fetch_add(&pool->next, 1);
        if (i >= pool->count) break;
        pool->fn(pool->arg, i);
    }
    return NULL;
}

int lp_oracle_parallel_for(int count, int threads, void (*fn)(void *, int), void *arg)
{
    if (threads <= 0) {
        long nc = sysconf(_SC_NPROCESSORS_ONLN);
        threads = nc > 0 ? (int)nc : 1;
    }
    if (threads > count) threads = count > 0 ? count : 1;
    pool_t pool;
    atomic_init(&pool.next, 0);
    pool.count = count;
    pool.fn = fn;
    pool.arg = arg;
    if (threads == 1) {
        worker(&pool);
        return 1;
    }
    pthread_t *tid = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
    int started = 0;
    for (int t = 0; t < threads; ++t)
        if (pthread_create(&tid[t], NULL, worker, &pool) == 0) started++; else break;
    if (started == 0) worker(&pool);
    for (int t = 0; t < started; ++t) pthread_join(tid[t], NULL);
    free(tid);
    return started > 0 ? started : 1;
}
