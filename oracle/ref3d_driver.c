/* oracle/ref3d_driver.c -- TEST INFRASTRUCTURE.  Drives the UNMODIFIED 3-D TTM functions of the
 * reference (imd_ttm_3D.c: do_DIFF :1603-1829, ttm_fill_ghost_layers :3314-3504,
 * ttm_create_mpi_datatypes :2972-3312) on a lattice read from a file, so that the 3-D stencil
 * restatement (ttm3d_np.py) and the CUDA kernel (imd_b200/csrc/ttm3d.cu) can be pinned to the
 * reference's own code.  imd_ttm_3D.c is an orphan source (not in the reference's Makefile) and
 * has no driver of its own; this file only supplies what init_ttm would have set up: the
 * globals, the two lattices and their pointer tables (same layout as :1066-1083).
 *
 * Linked as: this file (with MAIN defined: it owns the reference's globals) + imd_ttm_3D.o
 * (compiled from /root/reference, untouched) + shmpi, with --unresolved-symbols=ignore-all
 * because the rest of IMD is not linked: only the three functions above are ever called.
 *
 *   SHMPI_NP=1 ref3d_diff in.bin out.bin
 *
 * in.bin : int32[12] = { nx, ny, nz, fd_min_atoms, diff_substeps, laser_active, nsweeps,
 *                        pbc_x, pbc_y, pbc_z, 0, 0 }
 *          double[4] = { hx, hy, hz, tau }
 *          double[10][(nx+2)(ny+2)(nz+2)] = natoms, temp, fd_k, fd_g, Ce, md_temp, source, U, xi, dens
 *          (extended lattice, z fastest; both lattices start from the same values)
 * stdout : "ref3d: <nsweeps> sweeps of <cells> cells in <seconds> s" (the loop alone, for the CPU baseline)
 * out.bin: double[1] = Eabs_local; double[4][(nx+2)(ny+2)(nz+2)] = natoms, temp, U, xi of l1
 *          after nsweeps x { do_DIFF(tau); ttm_fill_ghost_layers(); }
 */
#define MAIN
#include "imd.h"

#include <stdint.h>

void error(char *msg) { fprintf(stderr, "ref3d_driver: reference error(): %s\n", msg); exit(2); }

static void rd(void *p, size_t n, FILE *f) { if (fread(p, 1, n, f) != n) { fprintf(stderr, "short read\n"); exit(1); } }

int main(int argc, char **argv)
{
  int32_t hd[12];
  double hv[4];
  int i, j, k, f, s;
  if (argc != 3) { fprintf(stderr, "usage: %s in.bin out.bin\n", argv[0]); return 1; }
  MPI_Init(&argc, &argv);
  MPI_Comm_rank(MPI_COMM_WORLD, &myid);
  MPI_Comm_size(MPI_COMM_WORLD, &num_cpus);
  if (num_cpus != 1) { fprintf(stderr, "ref3d_driver runs on one rank (self Sendrecv for the periodic ghost planes)\n"); return 1; }
  FILE *in = fopen(argv[1], "rb");
  if (!in) { perror(argv[1]); return 1; }
  rd(hd, sizeof hd, in);
  rd(hv, sizeof hv, in);
  local_fd_dim.x = hd[0] + 2; local_fd_dim.y = hd[1] + 2; local_fd_dim.z = hd[2] + 2;
  global_fd_dim.x = hd[0]; global_fd_dim.y = hd[1]; global_fd_dim.z = hd[2];
  fd_min_atoms = hd[3]; diff_substeps = hd[4]; laser_active = hd[5] != 0;
  pbc_dirs.x = hd[7]; pbc_dirs.y = hd[8]; pbc_dirs.z = hd[9];
  fd_h.x = hv[0]; fd_h.y = hv[1]; fd_h.z = hv[2];
  fd_vol = fd_h.x * fd_h.y * fd_h.z;                         /* imd_ttm_3D.c init_ttm */
  my_coord.x = my_coord.y = my_coord.z = 0;
  cpu_dim.x = cpu_dim.y = cpu_dim.z = 1;
  cpugrid = MPI_COMM_WORLD;
  nbwest = nbeast = nbnorth = nbsouth = nbup = nbdown = 0;
  steps = 1;
  Eabs_local = 0.0;
  const size_t n = (size_t) local_fd_dim.x * local_fd_dim.y * local_fd_dim.z;
  lattice1 = (ttm_Element *) calloc(n, sizeof(ttm_Element));
  lattice2 = (ttm_Element *) calloc(n, sizeof(ttm_Element));
  l1 = (ttm_Element ***) malloc(local_fd_dim.x * sizeof(ttm_Element **));
  l2 = (ttm_Element ***) malloc(local_fd_dim.x * sizeof(ttm_Element **));
  for (i = 0; i < local_fd_dim.x; i++) {
    l1[i] = (ttm_Element **) malloc(local_fd_dim.y * sizeof(ttm_Element *));
    l2[i] = (ttm_Element **) malloc(local_fd_dim.y * sizeof(ttm_Element *));
    for (j = 0; j < local_fd_dim.y; j++) {
      l1[i][j] = lattice1 + (size_t) i * local_fd_dim.y * local_fd_dim.z + (size_t) j * local_fd_dim.z;
      l2[i][j] = lattice2 + (size_t) i * local_fd_dim.y * local_fd_dim.z + (size_t) j * local_fd_dim.z;
    }
  }
  double *buf = (double *) malloc(n * sizeof(double));
  for (f = 0; f < 10; f++) {
    rd(buf, n * sizeof(double), in);
    for (size_t c = 0; c < n; c++) {
      ttm_Element *e1 = lattice1 + c, *e2 = lattice2 + c;
      const double v = buf[c];
      switch (f) {
        case 0: e1->natoms = e2->natoms = (int) v; break;
        case 1: e1->temp = e2->temp = v; break;
        case 2: e1->fd_k = e2->fd_k = v; break;
        case 3: e1->fd_g = e2->fd_g = v; break;
        case 4: e1->Ce = e2->Ce = v; break;
        case 5: e1->md_temp = e2->md_temp = v; break;
        case 6: e1->source = e2->source = v; break;
        case 7: e1->U = e2->U = v; break;
        case 8: e1->xi = e2->xi = v; break;
        case 9: e1->dens = e2->dens = v; break;
      }
    }
  }
  fclose(in);
  ttm_create_mpi_datatypes();
  const double t0 = MPI_Wtime();
  for (s = 0; s < hd[6]; s++) {
    do_DIFF(hv[3]);
    ttm_fill_ghost_layers();
  }
  const double t1 = MPI_Wtime();
  printf("ref3d: %d sweeps of %ld cells in %.6f s\n", (int) hd[6], (long) hd[0] * hd[1] * hd[2], t1 - t0);
  FILE *out = fopen(argv[2], "wb");
  if (!out) { perror(argv[2]); return 1; }
  fwrite(&Eabs_local, sizeof(double), 1, out);
  for (f = 0; f < 4; f++) {
    for (size_t c = 0; c < n; c++) {
      const ttm_Element *e = lattice1 == &l1[0][0][0] ? lattice1 + c : lattice2 + c;
      buf[c] = f == 0 ? (double) e->natoms : f == 1 ? e->temp : f == 2 ? e->U : e->xi;
    }
    fwrite(buf, sizeof(double), n, out);
  }
  fclose(out);
  (void) k;
  MPI_Finalize();
  return 0;
}
