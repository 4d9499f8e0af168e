/* See mpi.h.  Progress engine: every rank keeps FIFO queues of pending sends (per destination) and pending
 * receives (per source); one mailbox per directed rank pair carries one chunk (<= CAP bytes) at a time.
 * Blocking calls spin on progress(), so messages larger than a mailbox and sends nobody waits for
 * (the final gather, main.c:341) complete as they would under a real MPI library. */
#define _GNU_SOURCE
#include "mpi.h"
#include <sched.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#define CAP ((size_t)32 << 20)
#define MAXP 16

typedef struct {
    volatile int full;              /* 0: free for the sender, 1: holds a chunk for the receiver */
    volatile int tag;
    volatile size_t total, offset, len;
    char pad[64];
} mbox_hdr;

typedef struct {
    volatile int bar_count, bar_sense;
    volatile int abort_flag;
    mbox_hdr hdr[MAXP][MAXP];       /* [src][dst] */
} shared_ctl;

struct adv_mpi_req {
    int is_send, peer, tag, done;
    char* buf;
    size_t bytes, offset;
    struct adv_mpi_req* next;
};

static shared_ctl* g_ctl;
static char* g_data;                /* P*P mailboxes of CAP bytes, touched lazily */
static int g_rank, g_np = 1, g_sense;
static struct adv_mpi_req *g_sendq[MAXP], *g_recvq[MAXP];
static pid_t g_kids[MAXP];

static char* mbox_data(int src, int dst) { return g_data + ((size_t)src * g_np + dst) * CAP; }

static void enqueue(struct adv_mpi_req** q, struct adv_mpi_req* r)
{
    r->next = NULL;
    while (*q) q = &(*q)->next;
    *q = r;
}

/* watchdog: a blocking call that makes no progress for ADV_MPI_TIMEOUT_S seconds (default 120) aborts every
 * rank with a message instead of hanging the box */
static double g_block_t0;
static const char* g_block_what;
static double now_s(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }
static void block_begin(const char* what) { g_block_t0 = now_s(); g_block_what = what; }
static void block_check(void)
{
    static double limit = -1;
    if (limit < 0) { const char* e = getenv("ADV_MPI_TIMEOUT_S"); limit = e ? atof(e) : 120.0; }
    if (now_s() - g_block_t0 > limit) {
        fprintf(stderr, "mpi shim: rank %d stuck in %s for more than %g s -- aborting all ranks\n", g_rank, g_block_what, limit);
        g_ctl->abort_flag = 1;
        _exit(4);
    }
    sched_yield();
}

static void progress(void)
{
    if (g_ctl->abort_flag) _exit(3);
    for (int p = 0; p < g_np; p++) {
        struct adv_mpi_req* s = g_sendq[p];
        if (s) {
            mbox_hdr* h = &g_ctl->hdr[g_rank][p];
            if (!h->full) {
                size_t len = s->bytes - s->offset;
                if (len > CAP) len = CAP;
                memcpy(mbox_data(g_rank, p), s->buf + s->offset, len);
                h->tag = s->tag; h->total = s->bytes; h->offset = s->offset; h->len = len;
                __sync_synchronize();
                h->full = 1;
                s->offset += len;
                if (s->offset == s->bytes) { s->done = 1; g_sendq[p] = s->next; }
            }
        }
        struct adv_mpi_req* r = g_recvq[p];
        if (r) {
            mbox_hdr* h = &g_ctl->hdr[p][g_rank];
            if (h->full) {
                __sync_synchronize();
                if (h->tag != r->tag || h->total != r->bytes || h->offset != r->offset) {
                    fprintf(stderr, "mpi shim: rank %d expected tag %d (%zu B @%zu) from %d, mailbox holds tag %d (%zu B @%zu)\n",
                            g_rank, r->tag, r->bytes, r->offset, p, h->tag, (size_t)h->total, (size_t)h->offset);
                    g_ctl->abort_flag = 1;
                    _exit(3);
                }
                memcpy(r->buf + r->offset, mbox_data(p, g_rank), h->len);
                r->offset += h->len;
                __sync_synchronize();
                h->full = 0;
                if (r->offset == r->bytes) { r->done = 1; g_recvq[p] = r->next; }
            }
        }
    }
}

int MPI_Init(int* argc, char*** argv)
{
    (void)argc; (void)argv;
    const char* e = getenv("ADV_MPI_NP");
    g_np = e ? atoi(e) : 1;
    if (g_np < 1 || g_np > MAXP) { fprintf(stderr, "mpi shim: ADV_MPI_NP must be in [1,%d]\n", MAXP); return 1; }
    g_ctl = (shared_ctl*)mmap(NULL, sizeof(shared_ctl), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    g_data = (char*)mmap(NULL, (size_t)g_np * g_np * CAP, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (g_ctl == MAP_FAILED || g_data == MAP_FAILED) { perror("mpi shim: mmap"); return 1; }
    memset((void*)g_ctl, 0, sizeof(*g_ctl));
    g_rank = 0;
    fflush(stdout);
    for (int r = 1; r < g_np; r++) {
        pid_t pid = fork();
        if (pid < 0) { perror("mpi shim: fork"); return 1; }
        if (pid == 0) { g_rank = r; break; }
        g_kids[r] = pid;
    }
    return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm c, int* rank) { (void)c; *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm c, int* size) { (void)c; *size = g_np; return MPI_SUCCESS; }
int MPI_Errhandler_set(MPI_Comm c, MPI_Errhandler e) { (void)c; (void)e; return MPI_SUCCESS; }

double MPI_Wtime(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

int MPI_Barrier(MPI_Comm c)
{
    (void)c;
    g_sense = !g_sense;
    if (__sync_add_and_fetch(&g_ctl->bar_count, 1) == g_np) {
        g_ctl->bar_count = 0;
        __sync_synchronize();
        g_ctl->bar_sense = g_sense;
    } else {
        block_begin("MPI_Barrier");
        while (g_ctl->bar_sense != g_sense) { progress(); block_check(); }
    }
    return MPI_SUCCESS;
}

static struct adv_mpi_req* new_req(int is_send, void* buf, int count, MPI_Datatype t, int peer, int tag)
{
    struct adv_mpi_req* r = (struct adv_mpi_req*)calloc(1, sizeof(*r));
    r->is_send = is_send; r->peer = peer; r->tag = tag; r->buf = (char*)buf;
    r->bytes = (size_t)count * (size_t)t;      /* MPI_FLOAT = 4, MPI_DOUBLE = 8 */
    if (r->bytes == 0) r->done = 1;
    else enqueue(is_send ? &g_sendq[peer] : &g_recvq[peer], r);
    return r;
}

int MPI_Isend(const void* buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm c, MPI_Request* req)
{
    (void)c;
    *req = new_req(1, (void*)buf, count, t, dest, tag);
    progress();
    return MPI_SUCCESS;
}

int MPI_Irecv(void* buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request* req)
{
    (void)c;
    *req = new_req(0, buf, count, t, src, tag);
    progress();
    return MPI_SUCCESS;
}

int MPI_Wait(MPI_Request* req, MPI_Status* st)
{
    (void)st;
    struct adv_mpi_req* r = *req;
    block_begin(r->is_send ? "MPI_Wait(send)" : "MPI_Wait/MPI_Recv(recv)");
    while (!r->done) { progress(); if (!r->done) block_check(); }
    free(r);
    *req = NULL;
    return MPI_SUCCESS;
}

int MPI_Recv(void* buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status* st)
{
    MPI_Request r;
    MPI_Irecv(buf, count, t, src, tag, c, &r);
    return MPI_Wait(&r, st);
}

int MPI_Finalize(void)
{
    /* complete sends nobody waited for, then leave together */
    block_begin("MPI_Finalize (pending sends)");
    for (;;) {
        int pending = 0;
        for (int p = 0; p < g_np; p++) if (g_sendq[p]) pending = 1;
        if (!pending) break;
        progress();
        block_check();
    }
    MPI_Barrier(MPI_COMM_WORLD);
    if (g_rank == 0)
        for (int r = 1; r < g_np; r++) { int status; waitpid(g_kids[r], &status, 0); }
    return MPI_SUCCESS;
}
