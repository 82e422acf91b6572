/* oracle/nn_probe.c -- TEST INFRASTRUCTURE: probe of the reference's third-party
 * interpolation library "nn" (Pavel Sakov, v1.86.2; shipped only as
 * /root/reference/nn_interpol/libnn.a).  The reference builds a Delaunay triangulation of
 * every EOS table and interpolates linearly on it (imd_interpol.c:1263-1267,
 * imd_ttm.c:2517,2568).  This tool does the same calls on a table file and writes
 *   (a) which diagonal the triangulation put into every rectangle of the (rectilinear)
 *       point grid -- the device path takes that as an input (DESIGN.md "EOS tables");
 *   (b) lpi_interpolate_point at probe points, as golden vectors.
 *
 *   nn_probe <table.txt> <probes.bin | -> <out.bin>
 * table.txt : format of nn_read_table (is js / xmin xmax ymin ymax / is*js lines x y z)
 * probes.bin: n doubles pairs (x,y), raw; "-" for none
 * out.bin   : int32 is, js, nprobe, pad ; uint8 diag[(is-1)*(js-1)] padded to 8 ; double z[nprobe]
 *             diag 0: rectangle (i,j) split by (i,j)-(i+1,j+1); 1: by (i+1,j)-(i,j+1); 255: neither
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "nn.h"
#include "delaunay.h"

int main(int argc, char **argv)
{
  if (argc != 4) { fprintf(stderr, "usage: nn_probe table.txt probes.bin|- out.bin\n"); return 2; }
  FILE *f = fopen(argv[1], "r");
  if (!f) { perror(argv[1]); return 2; }
  int is, js;
  double xmin, xmax, ymin, ymax;
  if (fscanf(f, "%d %d", &is, &js) != 2 || fscanf(f, "%lf %lf %lf %lf", &xmin, &xmax, &ymin, &ymax) != 4) return 2;
  int np = is * js;
  point *pts = malloc(np * sizeof(point));
  for (int i = 0; i < np; i++)
    if (fscanf(f, "%lf %lf %lf", &pts[i].x, &pts[i].y, &pts[i].z) != 3) { fprintf(stderr, "short table\n"); return 2; }
  fclose(f);

  delaunay *d = delaunay_build(np, pts, 0, NULL, 0, NULL);
  lpi *l = lpi_build(d);

  unsigned char *diag = malloc((size_t) (is - 1) * (js - 1));
  memset(diag, 255, (size_t) (is - 1) * (js - 1));
  for (int t = 0; t < d->ntriangles; t++) {
    int *v = d->triangles[t].vids;
    int imin = is, jmin = js, has[2][2] = {{0, 0}, {0, 0}};
    for (int k = 0; k < 3; k++) { int i = v[k] / js, j = v[k] % js; if (i < imin) imin = i; if (j < jmin) jmin = j; }
    int ok = 1;
    for (int k = 0; k < 3; k++) {
      int di = v[k] / js - imin, dj = v[k] % js - jmin;
      if (di > 1 || dj > 1) ok = 0; else has[di][dj] = 1;
    }
    if (!ok || imin >= is - 1 || jmin >= js - 1) continue;
    /* a triangle containing both (0,0) and (1,1) lies on the main diagonal */
    unsigned char val = (has[0][0] && has[1][1]) ? 0 : 1;
    diag[(size_t) imin * (js - 1) + jmin] = val;
  }

  int nprobe = 0;
  double *pr = NULL;
  if (strcmp(argv[2], "-") != 0) {
    FILE *g = fopen(argv[2], "rb");
    if (!g) { perror(argv[2]); return 2; }
    fseek(g, 0, SEEK_END); long sz = ftell(g); fseek(g, 0, SEEK_SET);
    nprobe = (int) (sz / (2 * sizeof(double)));
    pr = malloc(sz);
    if (fread(pr, 2 * sizeof(double), nprobe, g) != (size_t) nprobe) return 2;
    fclose(g);
  }
  double *z = malloc((nprobe + 1) * sizeof(double));
  for (int i = 0; i < nprobe; i++) {
    point p; p.x = pr[2 * i]; p.y = pr[2 * i + 1]; p.z = 0;
    lpi_interpolate_point(l, &p);
    z[i] = p.z;
  }
  FILE *o = fopen(argv[3], "wb");
  if (!o) { perror(argv[3]); return 2; }
  int hdr[4] = {is, js, nprobe, 0};
  fwrite(hdr, sizeof(int), 4, o);
  size_t nd = (size_t) (is - 1) * (js - 1), pad = (8 - nd % 8) % 8;
  fwrite(diag, 1, nd, o);
  unsigned char zero[8] = {0};
  fwrite(zero, 1, pad, o);
  fwrite(z, sizeof(double), nprobe, o);
  fclose(o);
  return 0;
}
