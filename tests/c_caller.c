/* A plain C99 caller of libtfsb200: what a JNI / cgo / Panama binding does, without any of them.
 *
 *   c_caller <graph.pb> <n_rows>
 *
 * graph.pb holds  z = x + 3  on a double column "x" (the README example of the reference, README.md:56-87).
 * The program imports the graph, asks for the node summaries (TensorFlowOps.analyzeGraphTF), resolves the output
 * shape for n_rows rows, runs the block from pinned host buffers and checks the result; it also provokes one
 * validation error to show the status / tfs_last_error convention.  Exit code 0 = all good.
 * Test infrastructure (tests/test_c_caller.py builds it with gcc -std=c99 and links it against the library). */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "tfs_b200.h"

#define CHECK(call)                                                              \
  do {                                                                           \
    int status_ = (call);                                                        \
    if (status_ != TFS_OK) {                                                     \
      fprintf(stderr, "%s -> %d: %s\n", #call, status_, tfs_last_error());       \
      return 2;                                                                  \
    }                                                                            \
  } while (0)

int main(int argc, char** argv) {
  if (argc != 3) {
    fprintf(stderr, "usage: %s graph.pb n_rows\n", argv[0]);
    return 64;
  }
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 65;
  static unsigned char graph_bytes[1 << 16];
  const size_t len = fread(graph_bytes, 1, sizeof graph_bytes, f);
  fclose(f);
  const long long n = atoll(argv[2]);

  tfs_graph* g = NULL;
  CHECK(tfs_graph_import(graph_bytes, len, &g));
  tfs_node_info info;
  CHECK(tfs_graph_node_info(g, "x", &info));
  if (!info.is_placeholder || info.dtype != TFS_DOUBLE) return 3;
  CHECK(tfs_graph_node_info(g, "z", &info));
  if (info.is_placeholder || info.dtype != TFS_DOUBLE || info.rank != 1 || info.dims[0] != -1) return 4;

  if (tfs_device_count() < 1) {
    printf("no device: import and analysis only\n");
    tfs_graph_free(g);
    return 0;
  }

  tfs_session* s = NULL;
  CHECK(tfs_session_create(g, 0, &s));

  double *x = NULL, *z = NULL;
  CHECK(tfs_host_alloc((size_t)n * sizeof(double), (void**)&x));
  CHECK(tfs_host_alloc((size_t)n * sizeof(double), (void**)&z));
  for (long long i = 0; i < n; ++i) x[i] = (double)i;

  tfs_tensor feed, fetch;
  memset(&feed, 0, sizeof feed);
  memset(&fetch, 0, sizeof fetch);
  feed.name = "x";
  feed.dtype = TFS_DOUBLE;
  feed.rank = 1;
  feed.dims[0] = n;
  feed.data = x;
  fetch.name = "z";
  CHECK(tfs_session_resolve(s, 1, &feed, 1, &fetch));
  if (fetch.dtype != TFS_DOUBLE || fetch.rank != 1 || fetch.dims[0] != n) return 5;
  fetch.data = z;
  CHECK(tfs_session_run(s, 1, &feed, 1, &fetch));
  for (long long i = 0; i < n; ++i)
    if (z[i] != (double)i + 3.0) {
      fprintf(stderr, "row %lld: got %g\n", i, z[i]);
      return 6;
    }

  /* the reference refuses implicit casts (impl/DebugRowOps.scala:328-333): so does the library, by status code */
  feed.dtype = TFS_FLOAT;
  if (tfs_session_run(s, 1, &feed, 1, &fetch) != TFS_ERR_INVALID) return 7;
  if (strstr(tfs_last_error(), "not compatible with the data type") == NULL) return 8;

  printf("ok: %lld rows, %lld kernel launches\n", n, (long long)tfs_session_launch_count(s));
  tfs_host_free(x);
  tfs_host_free(z);
  tfs_session_free(s);
  tfs_graph_free(g);
  return 0;
}
