/* oracle/mpi_stub/mpi.h -- TEST INFRASTRUCTURE.  Single-rank MPI shim (ours,
 * not MPI's) so that the reference sources, which #include <mpi.h>
 * (engine.hpp:6, engine.cpp:2), compile unmodified in a container without MPI.
 * Only the symbols the reference names are provided; the fork's MPI epoch
 * driver (engine.cpp:67-386) is never called by the oracle (it does not
 * terminate at one rank). */
#ifndef HEAT_ORACLE_MPI_STUB_H
#define HEAT_ORACLE_MPI_STUB_H
#include <string.h>
#include <stdint.h>
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
#define MPI_COMM_WORLD 0
#define MPI_SUM 0
#define MPI_FLOAT 4
#define MPI_DOUBLE 8
#define MPI_UINT64_T 9
#define MPI_UINT32_T 5
static inline int heat_stub_mpi_size(MPI_Datatype t) { return t == MPI_FLOAT || t == MPI_UINT32_T ? 4 : 8; }
static inline int MPI_Comm_rank(MPI_Comm c, int *r) { (void)c; *r = 0; return 0; }
static inline int MPI_Comm_size(MPI_Comm c, int *s) { (void)c; *s = 1; return 0; }
static inline int MPI_Bcast(void *b, int n, MPI_Datatype t, int root, MPI_Comm c)
{ (void)b; (void)n; (void)t; (void)root; (void)c; return 0; }
static inline int MPI_Allreduce(const void *s, void *r, int n, MPI_Datatype t, MPI_Op op, MPI_Comm c)
{ (void)op; (void)c; memcpy(r, s, (size_t)n * heat_stub_mpi_size(t)); return 0; }
#endif
