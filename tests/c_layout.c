/* Prints the layout of the structs that cross the C ABI.  A Panama MemoryLayout (or a JNI glue struct) is written
 * against these numbers; tests/test_c_caller.py pins them. */
#include <stddef.h>
#include <stdio.h>

#include "tfs_b200.h"

int main(void) {
  printf("tfs_tensor %zu name %zu dtype %zu rank %zu dims %zu data %zu\n", sizeof(tfs_tensor), offsetof(tfs_tensor, name),
         offsetof(tfs_tensor, dtype), offsetof(tfs_tensor, rank), offsetof(tfs_tensor, dims), offsetof(tfs_tensor, data));
  printf("tfs_node_info %zu dtype %zu rank %zu dims %zu is_placeholder %zu n_inputs %zu\n", sizeof(tfs_node_info),
         offsetof(tfs_node_info, dtype), offsetof(tfs_node_info, rank), offsetof(tfs_node_info, dims),
         offsetof(tfs_node_info, is_placeholder), offsetof(tfs_node_info, n_inputs));
  printf("tfs_shape_hint %zu name %zu rank %zu dims %zu\n", sizeof(tfs_shape_hint), offsetof(tfs_shape_hint, name),
         offsetof(tfs_shape_hint, rank), offsetof(tfs_shape_hint, dims));
  printf("tfs_node_summary %zu name %zu dtype %zu rank %zu dims %zu is_placeholder %zu is_input %zu is_output %zu\n",
         sizeof(tfs_node_summary), offsetof(tfs_node_summary, name), offsetof(tfs_node_summary, dtype),
         offsetof(tfs_node_summary, rank), offsetof(tfs_node_summary, dims), offsetof(tfs_node_summary, is_placeholder),
         offsetof(tfs_node_summary, is_input), offsetof(tfs_node_summary, is_output));
  return 0;
}
