/* Thread-per-rank implementation of baseline/mpi_shim/mpi.h.
 *   ./prog -n N <program args...>      (N defaults to $SHIM_NP or 1)
 * Sends are eager (packed copy queued on the destination); a receive is matched and
 * unpacked when it is waited on, which makes the reference's racy corner pixels
 * deterministic.  Messages between one (source, destination) pair are
 * delivered in order; the reference only uses tag 0. */
#define _GNU_SOURCE
#include "mpi.h"

#include <fcntl.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

struct shim_dtype { int count, blocklen, stride, elem; };   /* count blocks of blocklen elems, stride elems apart */
struct shim_dtype shim_byte = {1, 1, 1, 1}, shim_int = {1, 1, 1, 4}, shim_double = {1, 1, 1, 8};

struct msg { struct msg *next; size_t n; unsigned char data[]; };
struct shim_req { int is_recv, peer; void *buf; int count; struct shim_dtype t; };
struct shim_file { int fd; MPI_Offset pos; };

static int g_n = 1;
static pthread_barrier_t g_bar;
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t g_cv = PTHREAD_COND_INITIALIZER;
static struct msg **g_head, **g_tail;      /* [src * n + dst] FIFO */
static void *g_bcast;
static __thread int t_rank;

int ref_main(int argc, char **argv);

static size_t packed_size(int count, const struct shim_dtype *t) { return (size_t)count * t->count * t->blocklen * t->elem; }

static void pack(unsigned char *dst, const void *src, int count, const struct shim_dtype *t) {
    const unsigned char *s = src;
    size_t blk = (size_t)t->blocklen * t->elem, ext = (size_t)t->stride * t->elem;
    for (int c = 0; c < count; ++c)
        for (int i = 0; i < t->count; ++i, dst += blk) memcpy(dst, s + ((size_t)c * t->count + i) * ext, blk);
}
static void unpack(void *dstv, const unsigned char *src, int count, const struct shim_dtype *t) {
    unsigned char *d = dstv;
    size_t blk = (size_t)t->blocklen * t->elem, ext = (size_t)t->stride * t->elem;
    for (int c = 0; c < count; ++c)
        for (int i = 0; i < t->count; ++i, src += blk) memcpy(d + ((size_t)c * t->count + i) * ext, src, blk);
}

static void enqueue(int src, int dst, const void *buf, int count, const struct shim_dtype *t) {
    size_t n = packed_size(count, t);
    struct msg *m = malloc(sizeof *m + n);
    m->next = NULL; m->n = n;
    pack(m->data, buf, count, t);
    pthread_mutex_lock(&g_mu);
    int q = src * g_n + dst;
    if (g_tail[q]) g_tail[q]->next = m; else g_head[q] = m;
    g_tail[q] = m;
    pthread_cond_broadcast(&g_cv);
    pthread_mutex_unlock(&g_mu);
}
static void dequeue(int src, int dst, void *buf, int count, const struct shim_dtype *t) {
    pthread_mutex_lock(&g_mu);
    int q = src * g_n + dst;
    while (!g_head[q]) pthread_cond_wait(&g_cv, &g_mu);
    struct msg *m = g_head[q];
    g_head[q] = m->next;
    if (!g_head[q]) g_tail[q] = NULL;
    pthread_mutex_unlock(&g_mu);
    unpack(buf, m->data, count, t);
    free(m);
}

int MPI_Init(int *argc, char ***argv) { (void)argc; (void)argv; return 0; }
int MPI_Finalize(void) { return 0; }
int MPI_Abort(MPI_Comm c, int code) { (void)c; fflush(NULL); _exit(code); }
int MPI_Comm_size(MPI_Comm c, int *s) { (void)c; *s = g_n; return 0; }
int MPI_Comm_rank(MPI_Comm c, int *r) { (void)c; *r = t_rank; return 0; }
int MPI_Barrier(MPI_Comm c) { (void)c; pthread_barrier_wait(&g_bar); return 0; }
double MPI_Wtime(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + ts.tv_nsec * 1e-9; }

int MPI_Bcast(void *buf, int count, MPI_Datatype t, int root, MPI_Comm c) {
    (void)c;
    if (t_rank == root) g_bcast = buf;
    pthread_barrier_wait(&g_bar);
    if (t_rank != root) memcpy(buf, g_bcast, packed_size(count, t));
    pthread_barrier_wait(&g_bar);
    return 0;
}

int MPI_Type_vector(int count, int blocklen, int stride, MPI_Datatype old, MPI_Datatype *nt) {
    *nt = malloc(sizeof **nt);
    (*nt)->count = count; (*nt)->blocklen = blocklen; (*nt)->stride = stride; (*nt)->elem = old->elem;
    return 0;
}
int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype *nt) { return MPI_Type_vector(1, count, count, old, nt); }
int MPI_Type_commit(MPI_Datatype *t) { (void)t; return 0; }
int MPI_Type_free(MPI_Datatype *t) { free(*t); *t = NULL; return 0; }

int MPI_Isend(const void *buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm c, MPI_Request *req) {
    (void)tag; (void)c;
    enqueue(t_rank, dest, buf, count, t);
    *req = calloc(1, sizeof **req);
    return 0;
}
int MPI_Irecv(void *buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request *req) {
    (void)tag; (void)c;
    struct shim_req *r = calloc(1, sizeof *r);
    r->is_recv = 1; r->peer = src; r->buf = buf; r->count = count; r->t = *t;
    *req = r;
    return 0;
}
int MPI_Wait(MPI_Request *req, MPI_Status *st) {
    (void)st;
    struct shim_req *r = *req;
    if (r && r->is_recv) dequeue(r->peer, t_rank, r->buf, r->count, &r->t);
    free(r);
    *req = NULL;
    return 0;
}
int MPI_Send(const void *buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm c) {
    (void)tag; (void)c; enqueue(t_rank, dest, buf, count, t); return 0;
}
int MPI_Recv(void *buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status *st) {
    (void)tag; (void)c; (void)st; dequeue(src, t_rank, buf, count, t); return 0;
}

int MPI_File_open(MPI_Comm c, const char *name, int amode, MPI_Info info, MPI_File *fh) {
    (void)c; (void)info;
    int flags = (amode & MPI_MODE_WRONLY) ? O_WRONLY : O_RDONLY;
    if (amode & MPI_MODE_CREATE) flags |= O_CREAT;
    *fh = malloc(sizeof **fh);
    (*fh)->fd = open(name, flags, 0644);
    (*fh)->pos = 0;
    pthread_barrier_wait(&g_bar);   /* collective */
    return (*fh)->fd < 0;
}
int MPI_File_seek(MPI_File fh, MPI_Offset off, int whence) { (void)whence; fh->pos = off; return 0; }
int MPI_File_read(MPI_File fh, void *buf, int count, MPI_Datatype t, MPI_Status *st) {
    (void)st;
    size_t n = packed_size(count, t);
    ssize_t k = pread(fh->fd, buf, n, (off_t)fh->pos);
    if (k > 0) fh->pos += k;
    return k != (ssize_t)n;
}
int MPI_File_write(MPI_File fh, const void *buf, int count, MPI_Datatype t, MPI_Status *st) {
    (void)st;
    size_t n = packed_size(count, t);
    ssize_t k = pwrite(fh->fd, buf, n, (off_t)fh->pos);
    if (k > 0) fh->pos += k;
    return k != (ssize_t)n;
}
int MPI_File_close(MPI_File *fh) { close((*fh)->fd); free(*fh); *fh = NULL; return 0; }

struct targ { int rank, argc; char **argv; int rc; };
static void *rank_main(void *p) {
    struct targ *a = p;
    t_rank = a->rank;
    /* every rank gets its own argv copy, like under mpiexec */
    char **av = malloc((a->argc + 1) * sizeof *av);
    for (int i = 0; i < a->argc; ++i) av[i] = strdup(a->argv[i]);
    av[a->argc] = NULL;
    a->rc = ref_main(a->argc, av);
    return NULL;
}

#undef main
int main(int argc, char **argv) {
    const char *env = getenv("SHIM_NP");
    g_n = env ? atoi(env) : 1;
    int shift = 0;
    if (argc >= 3 && !strcmp(argv[1], "-n")) { g_n = atoi(argv[2]); shift = 2; }
    if (g_n < 1) g_n = 1;
    argv[shift] = argv[0];
    argc -= shift; argv += shift;
    g_head = calloc((size_t)g_n * g_n, sizeof *g_head);
    g_tail = calloc((size_t)g_n * g_n, sizeof *g_tail);
    pthread_barrier_init(&g_bar, NULL, (unsigned)g_n);
    pthread_t *th = malloc(sizeof *th * g_n);
    struct targ *ta = malloc(sizeof *ta * g_n);
    for (int i = 0; i < g_n; ++i) {
        ta[i].rank = i; ta[i].argc = argc; ta[i].argv = argv; ta[i].rc = 0;
        pthread_create(&th[i], NULL, rank_main, &ta[i]);
    }
    int rc = 0;
    for (int i = 0; i < g_n; ++i) { pthread_join(th[i], NULL); rc |= ta[i].rc; }
    return rc;
}
