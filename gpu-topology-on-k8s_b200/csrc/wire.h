// wire.h -- the extender's HTTP bodies as plain structs (internal; the C-ABI is gt_wire_* in gputopo.h)
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace gt {

struct ExtenderArgsWire {
    uint32_t gpu_count = 0;            // sum over containers of limits["aliyun.com/gpu"] (requests where no limit)
    std::string uid, name, ns;         // pod metadata
    bool has_names = false;            // NodeNames present (nodeCacheCapable: true, design.md:101)
    std::vector<std::string> names;
};

struct BindingArgsWire {
    std::string pod_name, pod_ns, pod_uid, node;
};

// GT_OK or GT_ERR_PARSE (malformed JSON, wrong types, fractional / negative GPU quantity)
int parse_extender_args(const char *json, size_t len, ExtenderArgsWire *out);
int parse_binding_args(const char *json, size_t len, BindingArgsWire *out);
void json_escape(const std::string &s, std::string *out);   // appends the quoted, escaped string

}  // namespace gt
