/* extender_demo.c -- the C-ABI used from plain C, the way a cgo shim would: load two nodes from the
 * `nvidia-smi topo -m` text of tests/golden, answer a `sort` body, `bind` the pod, delete it again.
 *
 *   gcc -std=c99 -Wall -Wextra -pedantic -Iinclude examples/extender_demo.c \
 *       -Lgpu-topology-on-k8s_b200 -lgputopo -Wl,-rpath,$PWD/gpu-topology-on-k8s_b200 -o /tmp/extender_demo
 *   /tmp/extender_demo tests/golden/f1_ring.topo.txt        (needs a CUDA device: there is no CPU fallback)
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "gputopo.h"

static int die(gt_ctx *ctx, const char *what, int rc)
{
    fprintf(stderr, "%s: %s (%d) %s\n", what, gt_strerror(rc), rc, ctx ? gt_last_error(ctx) : "");
    if (ctx) gt_destroy(ctx);
    return 1;
}

int main(int argc, char **argv)
{
    static char text[1 << 16], reply[1 << 12], annotations[256];
    uint8_t levels[2 * 256], n_gpus[2];
    uint16_t free_mask[2] = {0x00FF, 0x00F3};   /* node-b already runs a pod on GPUs 2 and 3 */
    uint32_t n = 0;
    size_t len, out_len = 0;
    gt_config cfg;
    gt_ctx *ctx = NULL;
    int rc;
    FILE *f;
    const char *sort_body =
        "{\"Pod\":{\"metadata\":{\"name\":\"train-0\",\"namespace\":\"default\",\"uid\":\"uid-1\"},"
        "\"spec\":{\"containers\":[{\"resources\":{\"limits\":{\"" GT_RESOURCE_NAME "\":\"2\"}}}]}},"
        "\"NodeNames\":[\"node-a\",\"node-b\"]}";
    const char *bind_body =
        "{\"PodName\":\"train-0\",\"PodNamespace\":\"default\",\"PodUID\":\"uid-1\",\"Node\":\"node-b\"}";

    if (argc < 2 || !(f = fopen(argv[1], "rb"))) {
        fprintf(stderr, "usage: %s <nvidia-smi topo -m text>\n", argv[0]);
        return 2;
    }
    len = fread(text, 1, sizeof text - 1, f);
    fclose(f);
    if ((rc = gt_parse_topo_matrix(text, len, levels, &n)) != GT_OK) return die(NULL, "gt_parse_topo_matrix", rc);
    memcpy(levels + 256, levels, 256);          /* two machines of the same type */
    n_gpus[0] = n_gpus[1] = (uint8_t)n;

    memset(&cfg, 0, sizeof cfg);
    cfg.struct_size = (uint32_t)sizeof cfg;
    cfg.max_nodes = 2;
    if ((rc = gt_create(&cfg, &ctx)) != GT_OK) return die(NULL, "gt_create", rc);
    if ((rc = gt_load_nodes(ctx, 0, 2, levels, n_gpus, NULL, NULL, free_mask)) != GT_OK) return die(ctx, "gt_load_nodes", rc);
    if ((rc = gt_set_node_names(ctx, "node-a\nnode-b", 13)) != GT_OK) return die(ctx, "gt_set_node_names", rc);

    if ((rc = gt_ext_sort_json(ctx, sort_body, strlen(sort_body), reply, sizeof reply, &out_len)) != GT_OK)
        return die(ctx, "gt_ext_sort_json", rc);
    printf("sort  -> %s\n", reply);
    if ((rc = gt_ext_bind_json(ctx, bind_body, strlen(bind_body), 1700000000, reply, sizeof reply, annotations,
                               sizeof annotations)) != GT_OK)
        return die(ctx, "gt_ext_bind_json", rc);
    printf("bind  -> %s\n%s", reply, annotations);
    if ((rc = gt_ext_pod_deleted(ctx, "uid-1")) != GT_OK) return die(ctx, "gt_ext_pod_deleted", rc);
    printf("pod deleted, GPUs released\n");
    gt_destroy(ctx);
    return 0;
}
