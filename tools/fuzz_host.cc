// fuzz_host.cc -- throws mutated and random bytes at the host-only parsers of the C-ABI under ASan/UBSan:
//   g++ -std=c++17 -O1 -g -fsanitize=address,undefined -fno-sanitize-recover=all -Iinclude tools/fuzz_host.cc \
//       gpu-topology-on-k8s_b200/csrc/{extender,wire,gaia_tree,host}.cc -o /tmp/fuzz_host && /tmp/fuzz_host 200000
// (these four files contain no CUDA; api.cu is not linked, so nothing here can reach a kernel)
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "gputopo.h"

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t rnd()
{
    rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17;
    return rng_state;
}

static std::string mutate(const std::string &seed)
{
    std::string s = seed;
    const int edits = 1 + (int)(rnd() % 6);
    static const char pool[] = "{}[]\",:\\ \n\tnullu0019Xx-+.eE_GPUNVSY";
    for (int e = 0; e < edits && !s.empty(); e++) {
        const size_t at = rnd() % s.size();
        switch (rnd() % 5) {
        case 0: s[at] = pool[rnd() % (sizeof pool - 1)]; break;
        case 1: s.erase(at, 1 + rnd() % 4); break;
        case 2: s.insert(at, 1, pool[rnd() % (sizeof pool - 1)]); break;
        case 3: s.insert(at, s.substr(rnd() % s.size(), rnd() % 16)); break;
        default: s[at] = (char)(rnd() & 0xFF); break;
        }
    }
    if (rnd() % 8 == 0) s.resize(rnd() % (s.size() + 1));
    return s;
}

int main(int argc, char **argv)
{
    const long iters = argc > 1 ? atol(argv[1]) : 100000;
    const std::vector<std::string> json = {
        "{\"Pod\":{\"metadata\":{\"name\":\"p\",\"namespace\":\"d\",\"uid\":\"u-1\"},\"spec\":{\"containers\":[{\"resources\":"
        "{\"limits\":{\"aliyun.com/gpu\":\"2\",\"cpu\":\"500m\"},\"requests\":{\"aliyun.com/gpu\":1}}},{\"resources\":null}]}},"
        "\"Nodes\":null,\"NodeNames\":[\"n1\",\"n\\u00e9\\ud83d\\ude00\",\"a\\\"b\"]}",
        "{\"PodName\":\"p\",\"PodNamespace\":\"d\",\"PodUID\":\"u-1\",\"Node\":\"n1\"}",
    };
    const std::vector<std::string> text = {
        "\tGPU0\tGPU1\tGPU2\tCPU Affinity\nGPU0\t X \tNV2\tSYS\t0-31\nGPU1\tNV2\t X \tPHB\t0-31\nGPU2\tSYS\tPHB\t X \t32-63\n\nLegend:\n",
        "GPU_SYS_0_1: Cross CPU socket\nGPU_NV2_0_2: Two NVLINK links\nGPU_PIX_1_2: Single PCI switch\n",
        "ALIYUN_COM_GPU_ASSIGNED: false\nALIYUN_COM_GPU_ASSUME_TIME: 1561717704\nALIYUN_COM_GPU_GROUP: 0,1,2,3\n",
        "0,1,2,15",
    };
    char a[64], b[64], c[64], d[64], big[4096];
    uint8_t levels[256], access[16];
    uint16_t sibling[16], mask;
    uint32_t n, cnt;
    int32_t nn;
    uint8_t ng;
    size_t len;
    long parsed = 0;
    for (long it = 0; it < iters; it++) {
        const std::string j = mutate(json[rnd() % json.size()]);
        const size_t cap = rnd() % 3 == 0 ? rnd() % 40 : sizeof big;      // small buffers too: GT_ERR_RANGE, not overflow
        parsed += gt_wire_parse_extender_args(j.data(), j.size(), &cnt, a, rnd() % sizeof a + 1, big, cap + 1 > sizeof big ? sizeof big : cap + 1, &nn) == GT_OK;
        parsed += gt_wire_parse_binding_args(j.data(), j.size(), a, sizeof a, b, rnd() % sizeof b + 1, c, sizeof c, d, sizeof d) == GT_OK;
        const std::string t = mutate(text[rnd() % text.size()]);
        parsed += gt_parse_topo_matrix(t.data(), t.size(), levels, &n) == GT_OK;
        parsed += gt_annotation_decode(t.data(), t.size(), levels, &n) == GT_OK;
        parsed += gt_group_decode(t.c_str(), &mask) == GT_OK;
        parsed += gt_plugin_allocate(t.c_str(), 1700000000, a, rnd() % sizeof a + 1, big, cap + 1 > sizeof big ? sizeof big : cap + 1) == GT_OK;
        // encoders with tight buffers
        for (int i = 0; i < 256; i++) levels[i] = (uint8_t)(rnd() & 15);
        gt_annotation_encode(levels, 1 + (uint32_t)(rnd() % 16), big, rnd() % sizeof big, &len);
        gt_group_encode((uint16_t)rnd(), a, rnd() % sizeof a);
        gt_bind_annotations((uint16_t)rnd(), (int64_t)rnd(), (int)(rnd() & 1), big, rnd() % 200);
        gt_wire_binding_result(t.c_str(), big, rnd() % 64 + 1);
        const uint8_t sc[3] = {0, 10, 7};
        gt_wire_host_priorities(t.c_str(), (uint32_t)(rnd() % 4), sc, big, rnd() % 128 + 1, &len);
        // trees: random structures, valid or not
        gt_tree tr;
        tr.n_nodes = (int32_t)(rnd() % 36) - 2;
        for (int i = 0; i < GT_MAX_TREE_NODES; i++) {
            tr.parent[i] = (int32_t)(rnd() % 40) - 4;
            tr.gpu_id[i] = rnd() % 3 ? (int32_t)(rnd() % 20) - 2 : -1;
            tr.access_cost[i] = (int32_t)(rnd() % 300) - 10;
            tr.level[i] = (int32_t)(rnd() % 20) - 2;
        }
        parsed += gt_tree_tables(&tr, levels, &ng, access, sibling) == GT_OK;
        int32_t types[256];
        for (int i = 0; i < 256; i++) types[i] = (int32_t)(rnd() % 16) - 1;
        gt_levels_from_p2plink(types, (uint32_t)(rnd() % 18), levels);
    }
    printf("fuzz_host: %ld iterations, %ld inputs accepted, no sanitizer report\n", iters, parsed);
    return 0;
}
