// extender.cc -- host-only text formats on either side of the hot path (SURVEY.md section 8f row 1 and the
// annotation half of row 4): the `nvidia-smi topo -m` matrix parser, the node-annotation codec
// (design.md:78-82), the pod annotations written by bind (design.md:223-234) and the device plugin's
// Allocate reply (design.md:236-246).  No subset scoring here.
#include "../../include/gputopo.h"

#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {

struct LinkName {
    int code;
    const char *abbr;
    const char *desc;   // P2PLinkTypeDesc column of design.md:33-44 (NVLink rows are blank there)
};
const LinkName kLinks[] = {
    {GT_LV_SYS, "SYS", "Cross CPU socket"},       {GT_LV_NODE, "NODE", "Same CPU socket"},
    {GT_LV_PHB, "PHB", "Host PCI bridge"},        {GT_LV_PXB, "PXB", "Multiple PCI switches"},
    {GT_LV_PIX, "PIX", "Single PCI switch"},      {GT_LV_PSB, "PSB", "Same board"},
    {GT_LV_NV1, "NV1", "Single NVLINK link"},     {GT_LV_NV2, "NV2", "Two NVLINK links"},
    {GT_LV_NV3, "NV3", "Three NVLINK links"},     {GT_LV_NV4, "NV4", "Four NVLINK links"},
    {GT_LV_NVS, "NVS", "NVSwitch"},
};

// token of the topo matrix or of an annotation key -> level code, -1 if unknown
int level_of(const std::string &tok)
{
    for (const LinkName &l : kLinks)
        if (tok == l.abbr) return l.code;
    if (tok.size() > 2 && tok[0] == 'N' && tok[1] == 'V') {   // NV# with # > 4: a switch-attached bonded set
        char *end = nullptr;
        const long c = strtol(tok.c_str() + 2, &end, 10);
        if (*end == 0 && c >= 1) return c == 1 ? GT_LV_NV1 : c == 2 ? GT_LV_NV2 : c == 3 ? GT_LV_NV3 : c == 4 ? GT_LV_NV4 : GT_LV_NVS;
    }
    return -1;
}

const LinkName *link_of(int code)
{
    for (const LinkName &l : kLinks)
        if (l.code == code) return &l;
    return nullptr;
}

// strips ANSI escape sequences (nvidia-smi underlines the header) and splits into lines of tokens
std::vector<std::vector<std::string>> tokenize(const char *text, size_t len)
{
    std::vector<std::vector<std::string>> lines(1);
    std::string cur;
    auto flush = [&]() {
        if (!cur.empty()) { lines.back().push_back(cur); cur.clear(); }
    };
    for (size_t i = 0; i < len; i++) {
        const unsigned char ch = (unsigned char)text[i];
        if (ch == 0x1b && i + 1 < len && text[i + 1] == '[') {   // ESC [ ... letter
            i += 2;
            while (i < len && !isalpha((unsigned char)text[i])) i++;
            continue;
        }
        if (ch == '\n') { flush(); lines.emplace_back(); }
        else if (isspace(ch)) flush();
        else cur.push_back((char)ch);
    }
    flush();
    return lines;
}

bool gpu_index(const std::string &tok, int *idx)   // "GPU12" -> 12
{
    if (tok.size() < 4 || tok.compare(0, 3, "GPU") != 0) return false;
    char *end = nullptr;
    const long v = strtol(tok.c_str() + 3, &end, 10);
    if (*end != 0 || v < 0) return false;
    *idx = (int)v;
    return true;
}

int put(char *out, size_t cap, size_t *pos, const std::string &s)
{
    if (*pos + s.size() + 1 > cap) return GT_ERR_RANGE;
    memcpy(out + *pos, s.data(), s.size());
    *pos += s.size();
    out[*pos] = 0;
    return GT_OK;
}

std::string ids_of_mask(uint16_t mask)   // "0,1,2,3" as in ALIYUN_COM_GPU_GROUP (design.md:231)
{
    std::string s;
    for (int g = 0; g < GT_MAX_GPUS; g++)
        if ((mask >> g) & 1) {
            if (!s.empty()) s += ',';
            s += std::to_string(g);
        }
    return s;
}

}  // namespace

extern "C" {

int gt_parse_topo_matrix(const char *text, size_t len, uint8_t out_levels[256], uint32_t *out_n)
{
    if (!text || !out_levels || !out_n) return GT_ERR_INVALID;
    memset(out_levels, 0, 256);
    *out_n = 0;
    const auto lines = tokenize(text, len);
    // header: the line whose first token is GPU0; its leading GPU<i> tokens are the GPU columns
    int n = 0;
    size_t row0 = 0;
    for (size_t li = 0; li < lines.size(); li++) {
        int idx;
        if (!lines[li].empty() && gpu_index(lines[li][0], &idx) && idx == 0 && lines[li].size() >= 2 &&
            gpu_index(lines[li][1], &idx)) {
            for (const std::string &t : lines[li]) {
                if (!gpu_index(t, &idx) || idx != n) break;
                n++;
            }
            row0 = li + 1;
            break;
        }
        if (lines[li].size() == 1 && gpu_index(lines[li][0], &idx) && idx == 0) {   // single-GPU box: header "GPU0"
            n = 1;
            row0 = li + 1;
            break;
        }
    }
    if (n < 1 || n > GT_MAX_GPUS) return GT_ERR_PARSE;
    int rows = 0;
    for (size_t li = row0; li < lines.size() && rows < n; li++) {
        int idx;
        if (lines[li].empty() || !gpu_index(lines[li][0], &idx)) continue;
        if (idx != rows || (int)lines[li].size() < 1 + n) return GT_ERR_PARSE;
        for (int j = 0; j < n; j++) {
            const std::string &t = lines[li][1 + j];
            if (j == idx) {
                if (t != "X") return GT_ERR_PARSE;
                continue;
            }
            const int code = level_of(t);
            if (code < 0) return GT_ERR_PARSE;
            out_levels[idx * 16 + j] = (uint8_t)code;
        }
        rows++;
    }
    if (rows != n) return GT_ERR_PARSE;
    *out_n = (uint32_t)n;
    return GT_OK;
}

int gt_level_from_p2plink(int32_t p2p_link_type)
{
    // nvml.P2PLinkType of the Go bindings the design links to (design.md:50-52), in the order of the table of
    // design.md:33-44 with P2PLinkUnknown = 0 in front; five and six NVLink links came later in the same enum
    static const int8_t map[13] = {GT_LV_SYS /* unknown: assume the farthest */, GT_LV_SYS, GT_LV_NODE, GT_LV_PHB,
                                   GT_LV_PXB, GT_LV_PIX, GT_LV_PSB, GT_LV_NV1, GT_LV_NV2, GT_LV_NV3, GT_LV_NV4,
                                   GT_LV_NVS, GT_LV_NVS};
    if (p2p_link_type < 0 || p2p_link_type > 12) return -1;
    return map[p2p_link_type];
}

int gt_levels_from_p2plink(const int32_t *types, uint32_t n, uint8_t out_levels[256])
{
    if (!types || !out_levels || n < 1 || n > GT_MAX_GPUS) return GT_ERR_INVALID;
    memset(out_levels, 0, 256);
    for (uint32_t i = 0; i < n; i++)
        for (uint32_t j = 0; j < n; j++) {
            if (i == j) continue;                       // "gpu to itself is not recorded" design.md:19,73
            const int code = gt_level_from_p2plink(types[i * n + j]);
            if (code < 0) return GT_ERR_INVALID;
            out_levels[i * 16 + j] = (uint8_t)code;
        }
    return GT_OK;
}

int gt_annotation_encode(const uint8_t levels[256], uint32_t n_gpus, char *out, size_t cap, size_t *out_len)
{
    if (!levels || !out || cap == 0 || n_gpus > GT_MAX_GPUS) return GT_ERR_INVALID;
    size_t pos = 0;
    out[0] = 0;
    for (uint32_t i = 0; i < n_gpus; i++)
        for (uint32_t j = i + 1; j < n_gpus; j++) {
            const LinkName *l = link_of(levels[i * 16 + j]);
            if (!l) return GT_ERR_INVALID;
            // design.md:80: "GPU_SYS_0_1: Cross CPU socket"
            const std::string line = std::string("GPU_") + l->abbr + "_" + std::to_string(i) + "_" + std::to_string(j) +
                                     ": " + l->desc + "\n";
            const int rc = put(out, cap, &pos, line);
            if (rc != GT_OK) return rc;
        }
    if (out_len) *out_len = pos;
    return GT_OK;
}

int gt_annotation_decode(const char *text, size_t len, uint8_t out_levels[256], uint32_t *out_n)
{
    if (!text || !out_levels || !out_n) return GT_ERR_INVALID;
    memset(out_levels, 0, 256);
    uint32_t n = 0;
    size_t i = 0;
    while (i < len) {
        size_t e = i;
        while (e < len && text[e] != '\n') e++;
        std::string line(text + i, e - i);
        i = e + 1;
        const size_t colon = line.find(':');
        std::string key = line.substr(0, colon);
        while (!key.empty() && isspace((unsigned char)key.back())) key.pop_back();
        size_t b = 0;
        while (b < key.size() && (isspace((unsigned char)key[b]) || key[b] == '"')) b++;
        key = key.substr(b);
        while (!key.empty() && key.back() == '"') key.pop_back();
        if (key.compare(0, 4, "GPU_") != 0) continue;   // other annotations of the node are not ours
        // GPU_<ABBR>_<i>_<j>
        const size_t p1 = key.find('_', 4);
        if (p1 == std::string::npos) return GT_ERR_PARSE;
        const size_t p2 = key.find('_', p1 + 1);
        if (p2 == std::string::npos) return GT_ERR_PARSE;
        const int code = level_of(key.substr(4, p1 - 4));
        char *end = nullptr;
        const std::string si = key.substr(p1 + 1, p2 - p1 - 1), sj = key.substr(p2 + 1);
        const long gi = strtol(si.c_str(), &end, 10);
        if (si.empty() || *end != 0) return GT_ERR_PARSE;
        const long gj = strtol(sj.c_str(), &end, 10);
        if (sj.empty() || *end != 0) return GT_ERR_PARSE;
        if (code < 0 || gi < 0 || gj < 0 || gi >= GT_MAX_GPUS || gj >= GT_MAX_GPUS || gi == gj) return GT_ERR_PARSE;
        out_levels[gi * 16 + gj] = (uint8_t)code;
        out_levels[gj * 16 + gi] = (uint8_t)code;
        if ((uint32_t)gi + 1 > n) n = (uint32_t)gi + 1;
        if ((uint32_t)gj + 1 > n) n = (uint32_t)gj + 1;
    }
    // every pair below n must have been named (no self entries: design.md:19)
    for (uint32_t a = 0; a < n; a++)
        for (uint32_t b2 = a + 1; b2 < n; b2++)
            if (out_levels[a * 16 + b2] == 0) return GT_ERR_PARSE;
    *out_n = n;
    return GT_OK;
}

int gt_group_encode(uint16_t mask, char *out, size_t cap)
{
    if (!out || cap == 0) return GT_ERR_INVALID;
    size_t pos = 0;
    out[0] = 0;
    return put(out, cap, &pos, ids_of_mask(mask));
}

int gt_group_decode(const char *text, uint16_t *out_mask)
{
    if (!text || !out_mask) return GT_ERR_INVALID;
    uint16_t m = 0;
    const char *p = text;
    while (*p) {
        while (*p && (isspace((unsigned char)*p) || *p == ',' || *p == '"')) p++;
        if (!*p) break;
        if (!isdigit((unsigned char)*p)) return GT_ERR_PARSE;
        char *end = nullptr;
        const long v = strtol(p, &end, 10);
        if (v < 0 || v >= GT_MAX_GPUS || ((m >> v) & 1)) return GT_ERR_PARSE;
        m |= (uint16_t)(1u << v);
        p = end;
    }
    *out_mask = m;
    return GT_OK;
}

int gt_bind_annotations(uint16_t mask, int64_t assume_time, int assigned, char *out, size_t cap)
{
    if (!out || cap == 0) return GT_ERR_INVALID;
    // design.md:229-231 (bind) and design.md:244-245 (Allocate flips ASSIGNED to true)
    size_t pos = 0;
    out[0] = 0;
    int rc = put(out, cap, &pos, std::string("ALIYUN_COM_GPU_ASSIGNED: ") + (assigned ? "true" : "false") + "\n");
    if (rc == GT_OK) rc = put(out, cap, &pos, "ALIYUN_COM_GPU_ASSUME_TIME: " + std::to_string((long long)assume_time) + "\n");
    if (rc == GT_OK) rc = put(out, cap, &pos, "ALIYUN_COM_GPU_GROUP: " + ids_of_mask(mask) + "\n");
    return rc;
}

int gt_plugin_allocate(const char *pod_annotations, int64_t now, char *out_env, size_t env_cap, char *out_annotations,
                       size_t ann_cap)
{
    if (!pod_annotations || !out_env || env_cap == 0) return GT_ERR_INVALID;
    // design.md:239: read the device list bind wrote, hand it to nvidia-docker as NVIDIA_VISIBLE_DEVICES
    const char *key = "ALIYUN_COM_GPU_GROUP";
    const char *p = strstr(pod_annotations, key);
    if (!p) return GT_ERR_PARSE;
    p += strlen(key);
    while (*p && (*p == '"' || isspace((unsigned char)*p))) p++;
    if (*p != ':') return GT_ERR_PARSE;
    p++;
    std::string val;
    while (*p && *p != '\n') val.push_back(*p++);
    uint16_t mask = 0;
    int rc = gt_group_decode(val.c_str(), &mask);
    if (rc != GT_OK) return rc;
    if (mask == 0) return GT_ERR_PARSE;
    size_t pos = 0;
    out_env[0] = 0;
    rc = put(out_env, env_cap, &pos, "NVIDIA_VISIBLE_DEVICES=" + ids_of_mask(mask));
    if (rc != GT_OK) return rc;
    if (out_annotations && ann_cap) {   // design.md:244-245: the two fields Allocate updates
        size_t apos = 0;
        out_annotations[0] = 0;
        rc = put(out_annotations, ann_cap, &apos, "ALIYUN_COM_GPU_ASSIGNED: true\n");
        if (rc == GT_OK)
            rc = put(out_annotations, ann_cap, &apos, "ALIYUN_COM_GPU_ASSUME_TIME: " + std::to_string((long long)now) + "\n");
    }
    return rc;
}

}  // extern "C"
