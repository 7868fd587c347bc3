/*
 * fork()-based stand-in for the handful of MPI calls the reference lda-c makes
 * (oni-ml/oni-lda-c/lda-estimate.c:183-318, 441-459).  TEST INFRASTRUCTURE:
 * used only to build oracle/_ref/lda from the unmodified reference sources.
 *
 * The reference calls MPI_Init AFTER read_data (lda-estimate.c:438-441), so a
 * fork here hands every rank the already-parsed corpus, which is the same
 * state "mpiexec -n P" gives it (every rank parses the whole file).
 *
 *   ONI_MPI_STUB_NP        number of ranks (default 1)
 *   ONI_MPI_STUB_MAXCOUNT  largest Allreduce count in elements (default 1<<20)
 *
 * Also holds __wrap_time: the reference seeds its RNG from time()
 * (lda-estimate.c:414-416); linking with -Wl,--wrap=time lets a test fix the
 * seed through ONI_ORACLE_SEED without touching reference files.
 */
#define _GNU_SOURCE
#include "mpi.h"

#include <pthread.h>
#include <semaphore.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#define MAILBOX_SLOTS 4096

typedef struct {
    int source;
    int tag;
    double payload;
} stub_msg;

typedef struct {
    pthread_barrier_t barrier;
    sem_t mail_free;
    sem_t mail_used;
    pthread_mutex_t mail_lock;
    unsigned long mail_head;
    unsigned long mail_tail;
    stub_msg mail[MAILBOX_SLOTS];
} stub_shared;

static int g_rank = 0;
static int g_size = 1;
static size_t g_maxcount = 0;
static stub_shared* g_sh = NULL;
static double* g_slots = NULL; /* g_size * g_maxcount doubles (ints use the same storage) */
static pid_t* g_children = NULL;

static size_t type_size(MPI_Datatype t) { return t == MPI_INT ? sizeof(int) : sizeof(double); }

int MPI_Init(int* argc, char*** argv)
{
    (void)argc; (void)argv;
    const char* np = getenv("ONI_MPI_STUB_NP");
    g_size = np ? atoi(np) : 1;
    if (g_size < 1) g_size = 1;
    if (g_size == 1) return MPI_SUCCESS;

    const char* mc = getenv("ONI_MPI_STUB_MAXCOUNT");
    g_maxcount = mc ? (size_t)atol(mc) : ((size_t)1 << 20);

    g_sh = mmap(NULL, sizeof(stub_shared), PROT_READ | PROT_WRITE,
                MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    g_slots = mmap(NULL, sizeof(double) * g_maxcount * (size_t)g_size, PROT_READ | PROT_WRITE,
                   MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (g_sh == MAP_FAILED || g_slots == MAP_FAILED) { perror("mpi_stub mmap"); exit(97); }

    pthread_barrierattr_t ba;
    pthread_barrierattr_init(&ba);
    pthread_barrierattr_setpshared(&ba, PTHREAD_PROCESS_SHARED);
    pthread_barrier_init(&g_sh->barrier, &ba, (unsigned)g_size);
    pthread_mutexattr_t ma;
    pthread_mutexattr_init(&ma);
    pthread_mutexattr_setpshared(&ma, PTHREAD_PROCESS_SHARED);
    pthread_mutex_init(&g_sh->mail_lock, &ma);
    sem_init(&g_sh->mail_free, 1, MAILBOX_SLOTS);
    sem_init(&g_sh->mail_used, 1, 0);
    g_sh->mail_head = g_sh->mail_tail = 0;

    fflush(NULL); /* do not duplicate buffered stdio into the children */
    g_children = calloc((size_t)g_size, sizeof(pid_t));
    for (int r = 1; r < g_size; r++) {
        pid_t pid = fork();
        if (pid < 0) { perror("mpi_stub fork"); exit(97); }
        if (pid == 0) { g_rank = r; free(g_children); g_children = NULL; break; }
        g_children[r] = pid;
    }
    return MPI_SUCCESS;
}

int MPI_Finalize(void)
{
    if (g_size == 1) return MPI_SUCCESS;
    fflush(NULL);
    if (g_rank != 0) _exit(0);
    int bad = 0;
    for (int r = 1; r < g_size; r++) {
        int st = 0;
        waitpid(g_children[r], &st, 0);
        if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) bad = 1;
    }
    if (bad) { fprintf(stderr, "mpi_stub: a rank failed\n"); exit(98); }
    return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm comm, int* rank) { (void)comm; *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm comm, int* size) { (void)comm; *size = g_size; return MPI_SUCCESS; }

int MPI_Barrier(MPI_Comm comm)
{
    (void)comm;
    if (g_size > 1) pthread_barrier_wait(&g_sh->barrier);
    return MPI_SUCCESS;
}

/* SUM in rank order 0..P-1 on every rank, so all ranks get the same bits. */
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count,
                  MPI_Datatype type, MPI_Op op, MPI_Comm comm)
{
    (void)op; (void)comm;
    size_t nbytes = type_size(type) * (size_t)count;
    if (g_size == 1) { memmove(recvbuf, sendbuf, nbytes); return MPI_SUCCESS; }
    if ((size_t)count > g_maxcount) {
        fprintf(stderr, "mpi_stub: Allreduce count %d > ONI_MPI_STUB_MAXCOUNT %zu\n", count, g_maxcount);
        exit(96);
    }
    char* mine = (char*)(g_slots + (size_t)g_rank * g_maxcount);
    memcpy(mine, sendbuf, nbytes);
    pthread_barrier_wait(&g_sh->barrier);
    if (type == MPI_INT) {
        int* out = (int*)recvbuf;
        for (int i = 0; i < count; i++) {
            int acc = 0;
            for (int r = 0; r < g_size; r++) acc += ((const int*)(g_slots + (size_t)r * g_maxcount))[i];
            out[i] = acc;
        }
    } else {
        double* out = (double*)recvbuf;
        for (int i = 0; i < count; i++) {
            double acc = 0.0;
            for (int r = 0; r < g_size; r++) acc += (g_slots + (size_t)r * g_maxcount)[i];
            out[i] = acc;
        }
    }
    pthread_barrier_wait(&g_sh->barrier);
    return MPI_SUCCESS;
}

/* Only single-value messages to rank 0 are used (lda-estimate.c:287, 297). */
int MPI_Send(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm)
{
    (void)comm; (void)dest;
    if (count != 1 || type != MPI_DOUBLE || g_size == 1) { fprintf(stderr, "mpi_stub: unsupported Send\n"); exit(95); }
    sem_wait(&g_sh->mail_free);
    pthread_mutex_lock(&g_sh->mail_lock);
    stub_msg* m = &g_sh->mail[g_sh->mail_tail % MAILBOX_SLOTS];
    m->source = g_rank; m->tag = tag; m->payload = *(const double*)buf;
    g_sh->mail_tail++;
    pthread_mutex_unlock(&g_sh->mail_lock);
    sem_post(&g_sh->mail_used);
    return MPI_SUCCESS;
}

int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag,
             MPI_Comm comm, MPI_Status* status)
{
    (void)comm;
    if (count != 1 || type != MPI_DOUBLE || source != MPI_ANY_SOURCE || tag != MPI_ANY_TAG || g_size == 1) {
        fprintf(stderr, "mpi_stub: unsupported Recv\n"); exit(95);
    }
    sem_wait(&g_sh->mail_used);
    pthread_mutex_lock(&g_sh->mail_lock);
    stub_msg m = g_sh->mail[g_sh->mail_head % MAILBOX_SLOTS];
    g_sh->mail_head++;
    pthread_mutex_unlock(&g_sh->mail_lock);
    sem_post(&g_sh->mail_free);
    *(double*)buf = m.payload;
    if (status) { status->MPI_SOURCE = m.source; status->MPI_TAG = m.tag; status->MPI_ERROR = 0; }
    return MPI_SUCCESS;
}

/* -Wl,--wrap=time : fixed RNG seed for deterministic oracle runs. */
time_t __real_time(time_t* t);
time_t __wrap_time(time_t* t)
{
    const char* s = getenv("ONI_ORACLE_SEED");
    if (!s) return __real_time(t);
    time_t v = (time_t)atol(s);
    if (t) *t = v;
    return v;
}
