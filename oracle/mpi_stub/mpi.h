/* Single-rank MPI stand-in used ONLY to compile the unmodified reference
 * (/root/reference/d2q9-bgk.c) as the CPU oracle under oracle/_ref/.
 * TEST INFRASTRUCTURE - never linked into the product.
 *
 * World size is 1 and rank is 0, so:
 *   - MPI_Sendrecv to self is a memcpy (the reference's halo_exchange
 *     d2q9-bgk.c:353,379 then implements the periodic y wrap),
 *   - MPI_Reduce(SUM) over one rank is a copy (d2q9-bgk.c:445),
 *   - MPI_Send / MPI_Recv are never reached with size==1 (d2q9-bgk.c:410,442).
 * Also pulls in _mm_malloc/_mm_free, which the reference uses without
 * including a header for them (d2q9-bgk.c:545). */
#ifndef LBM_ORACLE_MPI_STUB_H
#define LBM_ORACLE_MPI_STUB_H
#include <string.h>
#include <stdlib.h>
#include <mm_malloc.h>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_FLOAT 4
#define MPI_SUM 1
#define MPI_SUCCESS 0

static inline int MPI_Init(int* argc, char*** argv) { (void)argc; (void)argv; return MPI_SUCCESS; }
static inline int MPI_Finalize(void) { return MPI_SUCCESS; }
static inline int MPI_Comm_size(MPI_Comm c, int* size) { (void)c; *size = 1; return MPI_SUCCESS; }
static inline int MPI_Comm_rank(MPI_Comm c, int* rank) { (void)c; *rank = 0; return MPI_SUCCESS; }

static inline int MPI_Sendrecv(const void* sbuf, int scount, MPI_Datatype st, int dest, int stag,
                               void* rbuf, int rcount, MPI_Datatype rt, int src, int rtag,
                               MPI_Comm c, MPI_Status* status) {
    (void)dest; (void)stag; (void)src; (void)rtag; (void)c; (void)status; (void)rt;
    size_t n = (size_t)(scount < rcount ? scount : rcount) * (size_t)st;
    memmove(rbuf, sbuf, n);
    return MPI_SUCCESS;
}
static inline int MPI_Reduce(const void* sbuf, void* rbuf, int count, MPI_Datatype t, MPI_Op op,
                             int root, MPI_Comm c) {
    (void)op; (void)root; (void)c;
    memmove(rbuf, sbuf, (size_t)count * (size_t)t);
    return MPI_SUCCESS;
}
static inline int MPI_Send(const void* b, int n, MPI_Datatype t, int d, int tag, MPI_Comm c) {
    (void)b; (void)n; (void)t; (void)d; (void)tag; (void)c; abort(); return 1;
}
static inline int MPI_Recv(void* b, int n, MPI_Datatype t, int s, int tag, MPI_Comm c, MPI_Status* st) {
    (void)b; (void)n; (void)t; (void)s; (void)tag; (void)c; (void)st; abort(); return 1;
}
#endif
