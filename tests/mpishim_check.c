/* Exercises oracle/ref_harness/mpishim with the reference drivers' own call pattern (main.c:208-346):
 * Isend right/left, Irecv, Wait x4 per stage, then Isend-to-0 nobody waits for + Recv gather on rank 0. */
#include <mpi.h>
#include <stdio.h>
#include <stdlib.h>

int main(int argc, char** argv)
{
    int rank, np;
    MPI_Init(&argc, &argv);
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &np);
    const int n = argc > 1 ? atoi(argv[1]) : 1000;          /* doubles per halo message */
    double *rs = malloc(sizeof(double) * n), *ls = malloc(sizeof(double) * n), *rr = malloc(sizeof(double) * n), *lr = malloc(sizeof(double) * n);
    MPI_Request q[4];
    int bad = 0;
    for (int stage = 0; stage < 5; stage++) {
        for (int i = 0; i < n; i++) { rs[i] = rank * 1000 + stage + 0.5 * i; ls[i] = -(rank * 1000 + stage + 0.25 * i); }
        if (rank < np - 1) MPI_Isend(rs, n, MPI_DOUBLE, rank + 1, 1, MPI_COMM_WORLD, &q[0]);
        if (rank > 0) MPI_Isend(ls, n, MPI_DOUBLE, rank - 1, 5, MPI_COMM_WORLD, &q[1]);
        if (rank < np - 1) MPI_Irecv(rr, n, MPI_DOUBLE, rank + 1, 5, MPI_COMM_WORLD, &q[2]);
        if (rank > 0) MPI_Irecv(lr, n, MPI_DOUBLE, rank - 1, 1, MPI_COMM_WORLD, &q[3]);
        if (rank < np - 1) MPI_Wait(&q[2], MPI_STATUS_IGNORE);
        if (rank > 0) MPI_Wait(&q[3], MPI_STATUS_IGNORE);
        if (rank < np - 1) MPI_Wait(&q[0], MPI_STATUS_IGNORE);
        if (rank > 0) MPI_Wait(&q[1], MPI_STATUS_IGNORE);
        for (int i = 0; i < n; i += 97) {
            if (rank < np - 1 && rr[i] != -((rank + 1) * 1000 + stage + 0.25 * i)) bad++;
            if (rank > 0 && lr[i] != (rank - 1) * 1000 + stage + 0.5 * i) bad++;
        }
        MPI_Barrier(MPI_COMM_WORLD);
    }
    /* gather */
    double* all = rank == 0 ? malloc(sizeof(double) * n) : NULL;
    MPI_Request g;
    for (int i = 0; i < n; i++) rs[i] = rank + 1e-3 * i;
    MPI_Isend(rs, n, MPI_DOUBLE, 0, 0, MPI_COMM_WORLD, &g);
    if (rank == 0) {
        MPI_Status st;
        for (int r = 0; r < np; r++) {
            MPI_Recv(all, n, MPI_DOUBLE, r, 0, MPI_COMM_WORLD, &st);
            for (int i = 0; i < n; i += 97) if (all[i] != r + 1e-3 * i) bad++;
        }
    }
    double t = MPI_Wtime();
    MPI_Finalize();
    printf("rank %d of %d: %s (t=%g)\n", rank, np, bad ? "BAD" : "SHIM_OK", t > 0 ? 1.0 : 0.0);
    return bad != 0;
}
