/* Plain-C consumer of include/stemgnn_b200.h: what a cgo / JNI / FFI binding sees.  Host-only calls
 * (version, workspace queries, argument errors), so it runs without a GPU:
 *   gcc -std=c99 -Wall -Wextra -pedantic -Iinclude examples/abi_probe.c -o build/abi_probe \
 *       -Lstem-gnn_b200 -lstemgnn_b200 -Wl,-rpath,$PWD/stem-gnn_b200 */
#include <stdio.h>
#include <string.h>

#include "stemgnn_b200.h"

int main(void) {
    size_t bytes = 0;
    stemgnn_aggregate_args agg;
    int rc;

    printf("abi %d\n", stemgnn_abi_version());
    rc = stemgnn_csr_build_workspace(1166243, 169343, &bytes);
    printf("csr_build_workspace rc=%d bytes=%lu\n", rc, (unsigned long)bytes);
    if (rc != STEMGNN_OK || bytes == 0) return 1;
    rc = stemgnn_aggregate_hub_workspace(1166243, &bytes);
    printf("aggregate_hub_workspace rc=%d bytes=%lu\n", rc, (unsigned long)bytes);
    if (rc != STEMGNN_OK || bytes != 4u * (4u + 1166243u / 256u + 1u)) return 2;

    memset(&agg, 0, sizeof agg);
    agg.n_rows = 4; agg.feat = 8; agg.num_edges = 3;       /* rowptr / out left NULL on purpose */
    rc = stemgnn_aggregate(&agg, NULL);
    printf("aggregate(NULL pointers) rc=%d msg=%s\n", rc, stemgnn_last_error());
    if (rc != STEMGNN_EINVAL || strstr(stemgnn_last_error(), "NULL") == NULL) return 3;
    printf("launches so far %lu\n", (unsigned long)stemgnn_launch_count());
    return stemgnn_launch_count() == 0 ? 0 : 4;
}
