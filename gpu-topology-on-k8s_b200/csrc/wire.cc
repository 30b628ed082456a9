// wire.cc -- host-only codec of the scheduler extender's HTTP bodies: the JSON of
// ExtenderArgs / HostPriorityList (`sort`, design.md:99,117-118) and ExtenderBindingArgs / ExtenderBindingResult
// (`bind`, design.md:100,119).  The Go types are upstream Kubernetes types the reference links to (design.md:121)
// and does not contain; field names are matched without regard to case, as encoding/json does, so both the
// tagged (v1: "pod", "nodenames", "host", "score") and the untagged spellings are read.  No scoring here.
#include "wire.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/gputopo.h"

namespace {

bool ieq(const std::string &a, const char *b)
{
    size_t i = 0;
    for (; i < a.size() && b[i]; i++) {
        char x = a[i], y = b[i];
        if (x >= 'A' && x <= 'Z') x = (char)(x - 'A' + 'a');
        if (y >= 'A' && y <= 'Z') y = (char)(y - 'A' + 'a');
        if (x != y) return false;
    }
    return i == a.size() && b[i] == 0;
}

void put_utf8(uint32_t cp, std::string *out)
{
    if (cp < 0x80) out->push_back((char)cp);
    else if (cp < 0x800) { out->push_back((char)(0xC0 | (cp >> 6))); out->push_back((char)(0x80 | (cp & 0x3F))); }
    else if (cp < 0x10000) {
        out->push_back((char)(0xE0 | (cp >> 12))); out->push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
        out->push_back((char)(0x80 | (cp & 0x3F)));
    } else {
        out->push_back((char)(0xF0 | (cp >> 18))); out->push_back((char)(0x80 | ((cp >> 12) & 0x3F)));
        out->push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out->push_back((char)(0x80 | (cp & 0x3F)));
    }
}

// Minimal recursive-descent JSON reader: values are either consumed by a callback or skipped.
class Reader {
public:
    Reader(const char *p, size_t n) : p_(p), end_(p + n) {}

    void ws() { while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) p_++; }
    bool peek(char c) { ws(); return p_ < end_ && *p_ == c; }
    bool eat(char c) { if (!peek(c)) return false; p_++; return true; }
    bool at_end() { ws(); return p_ == end_; }

    bool literal(const char *w)
    {
        ws();
        const size_t n = strlen(w);
        if ((size_t)(end_ - p_) < n || memcmp(p_, w, n) != 0) return false;
        p_ += n;
        return true;
    }

    bool hex4(uint32_t *v)
    {
        if (end_ - p_ < 4) return false;
        uint32_t r = 0;
        for (int i = 0; i < 4; i++) {
            const char ch = p_[i];
            uint32_t d;
            if (ch >= '0' && ch <= '9') d = (uint32_t)(ch - '0');
            else if (ch >= 'a' && ch <= 'f') d = (uint32_t)(ch - 'a' + 10);
            else if (ch >= 'A' && ch <= 'F') d = (uint32_t)(ch - 'A' + 10);
            else return false;
            r = r * 16 + d;
        }
        p_ += 4;
        *v = r;
        return true;
    }

    bool string(std::string *out)
    {
        if (!eat('"')) return false;
        if (out) out->clear();
        while (p_ < end_) {
            const unsigned char ch = (unsigned char)*p_++;
            if (ch == '"') return true;
            if (ch < 0x20) return false;
            if (ch != '\\') { if (out) out->push_back((char)ch); continue; }
            if (p_ >= end_) return false;
            const char e = *p_++;
            uint32_t cp = 0;
            switch (e) {
            case '"': cp = '"'; break;
            case '\\': cp = '\\'; break;
            case '/': cp = '/'; break;
            case 'b': cp = '\b'; break;
            case 'f': cp = '\f'; break;
            case 'n': cp = '\n'; break;
            case 'r': cp = '\r'; break;
            case 't': cp = '\t'; break;
            case 'u': {
                if (!hex4(&cp)) return false;
                if (cp >= 0xD800 && cp < 0xDC00) {   // surrogate pair
                    uint32_t lo = 0;
                    if (end_ - p_ < 2 || p_[0] != '\\' || p_[1] != 'u') return false;
                    p_ += 2;
                    if (!hex4(&lo) || lo < 0xDC00 || lo > 0xDFFF) return false;
                    cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                } else if (cp >= 0xDC00 && cp <= 0xDFFF) {
                    return false;
                }
                break;
            }
            default: return false;
            }
            if (out) put_utf8(cp, out);
        }
        return false;
    }

    // the raw text of a number
    bool number(std::string *raw)
    {
        ws();
        const char *s = p_;
        if (p_ < end_ && (*p_ == '-' || *p_ == '+')) p_++;
        bool digits = false;
        while (p_ < end_ && ((*p_ >= '0' && *p_ <= '9') || *p_ == '.' || *p_ == 'e' || *p_ == 'E' || *p_ == '+' || *p_ == '-')) {
            if (*p_ >= '0' && *p_ <= '9') digits = true;
            p_++;
        }
        if (!digits) return false;
        if (raw) raw->assign(s, (size_t)(p_ - s));
        return true;
    }

    template <class F>
    bool object(F on_key)   // on_key(key) consumes the value
    {
        if (!eat('{') || ++depth_ > 64) return false;
        if (eat('}')) { depth_--; return true; }
        std::string key;
        for (;;) {
            if (!string(&key) || !eat(':')) return false;
            if (!on_key(key)) return false;
            if (eat(',')) continue;
            if (eat('}')) { depth_--; return true; }
            return false;
        }
    }

    template <class F>
    bool array(F on_item)   // on_item() consumes one value
    {
        if (!eat('[') || ++depth_ > 64) return false;
        if (eat(']')) { depth_--; return true; }
        for (;;) {
            if (!on_item()) return false;
            if (eat(',')) continue;
            if (eat(']')) { depth_--; return true; }
            return false;
        }
    }

    bool skip()
    {
        ws();
        if (p_ >= end_) return false;
        switch (*p_) {
        case '{': return object([&](const std::string &) { return skip(); });
        case '[': return array([&]() { return skip(); });
        case '"': return string(nullptr);
        case 't': return literal("true");
        case 'f': return literal("false");
        case 'n': return literal("null");
        default: return number(nullptr);
        }
    }

    // a string value, or null (left empty)
    bool string_or_null(std::string *out)
    {
        if (peek('n')) { out->clear(); return literal("null"); }
        return string(out);
    }

private:
    const char *p_, *end_;
    int depth_ = 0;
};

// a Kubernetes quantity that must be a whole, non-negative number of GPUs: "2", 2, "+2", "2.0"
bool gpu_quantity(Reader &r, uint32_t *out)
{
    std::string raw;
    if (r.peek('"')) {
        if (!r.string(&raw)) return false;
    } else if (!r.number(&raw)) {
        return false;
    }
    size_t i = 0;
    if (i < raw.size() && raw[i] == '+') i++;
    uint64_t v = 0;
    size_t digits = 0;
    for (; i < raw.size() && raw[i] >= '0' && raw[i] <= '9'; i++, digits++) {
        v = v * 10 + (uint64_t)(raw[i] - '0');
        if (v > 1000000) return false;
    }
    if (!digits) return false;
    if (i < raw.size() && raw[i] == '.') {            // "2.0" is a whole number, "0.5" (Gaia's Fragment) is not served
        for (i++; i < raw.size(); i++)
            if (raw[i] != '0') return false;
    }
    if (i != raw.size()) return false;                // suffixes ("500m", "1k") and exponents are not GPU counts
    *out = (uint32_t)v;
    return true;
}

// {"limits": {...}, "requests": {...}} -> GPUs of one container
bool container_gpus(Reader &r, uint32_t *out)
{
    bool has_limit = false, has_request = false;
    uint32_t limit = 0, request = 0;
    auto list = [&](bool *has, uint32_t *val) {
        if (r.peek('n')) return r.literal("null");
        return r.object([&](const std::string &k) {
            if (k == GT_RESOURCE_NAME) { *has = true; return gpu_quantity(r, val); }
            return r.skip();
        });
    };
    const bool ok = r.peek('n') ? r.literal("null") : r.object([&](const std::string &k) {
        if (ieq(k, "limits")) return list(&has_limit, &limit);
        if (ieq(k, "requests")) return list(&has_request, &request);
        return r.skip();
    });
    if (!ok) return false;
    *out = has_limit ? limit : (has_request ? request : 0u);
    return true;
}

bool pod(Reader &r, gt::ExtenderArgsWire *out)
{
    if (r.peek('n')) return r.literal("null");
    return r.object([&](const std::string &k) {
        if (ieq(k, "metadata")) {
            if (r.peek('n')) return r.literal("null");
            return r.object([&](const std::string &m) {
                if (ieq(m, "uid")) return r.string_or_null(&out->uid);
                if (ieq(m, "name")) return r.string_or_null(&out->name);
                if (ieq(m, "namespace")) return r.string_or_null(&out->ns);
                return r.skip();
            });
        }
        if (ieq(k, "spec")) {
            if (r.peek('n')) return r.literal("null");
            return r.object([&](const std::string &s) {
                if (!ieq(s, "containers")) return r.skip();
                if (r.peek('n')) return r.literal("null");
                out->gpu_count = 0;
                return r.array([&]() {
                    return r.object([&](const std::string &c) {
                        if (!ieq(c, "resources")) return r.skip();
                        uint32_t g = 0;
                        if (!container_gpus(r, &g)) return false;
                        out->gpu_count += g;
                        return true;
                    });
                });
            });
        }
        return r.skip();
    });
}

int copy_out(const std::string &s, char *out, size_t cap)
{
    if (!out) return GT_OK;
    if (s.size() + 1 > cap) return GT_ERR_RANGE;
    memcpy(out, s.c_str(), s.size() + 1);
    return GT_OK;
}

}  // namespace

namespace gt {

void json_escape(const std::string &s, std::string *out)
{
    out->push_back('"');
    for (const char ch : s) {
        const unsigned char u = (unsigned char)ch;
        switch (ch) {
        case '"': *out += "\\\""; break;
        case '\\': *out += "\\\\"; break;
        case '\n': *out += "\\n"; break;
        case '\r': *out += "\\r"; break;
        case '\t': *out += "\\t"; break;
        default:
            if (u < 0x20) {
                char buf[8];
                snprintf(buf, sizeof buf, "\\u%04x", u);
                *out += buf;
            } else {
                out->push_back(ch);
            }
        }
    }
    out->push_back('"');
}

int parse_extender_args(const char *json, size_t len, ExtenderArgsWire *out)
{
    if (!json || !out) return GT_ERR_INVALID;
    *out = ExtenderArgsWire();
    Reader r(json, len);
    const bool ok = r.object([&](const std::string &k) {
        if (ieq(k, "pod")) return pod(r, out);
        if (ieq(k, "nodenames")) {
            out->names.clear();
            if (r.peek('n')) { out->has_names = false; return r.literal("null"); }
            out->has_names = true;
            return r.array([&]() {
                std::string n;
                if (!r.string(&n)) return false;
                out->names.push_back(n);
                return true;
            });
        }
        return r.skip();   // "nodes": full objects are not sent with nodeCacheCapable: true and not needed
    });
    return ok && r.at_end() ? GT_OK : GT_ERR_PARSE;
}

int parse_binding_args(const char *json, size_t len, BindingArgsWire *out)
{
    if (!json || !out) return GT_ERR_INVALID;
    *out = BindingArgsWire();
    Reader r(json, len);
    const bool ok = r.object([&](const std::string &k) {
        if (ieq(k, "podname")) return r.string_or_null(&out->pod_name);
        if (ieq(k, "podnamespace")) return r.string_or_null(&out->pod_ns);
        if (ieq(k, "poduid")) return r.string_or_null(&out->pod_uid);
        if (ieq(k, "node")) return r.string_or_null(&out->node);
        return r.skip();
    });
    return ok && r.at_end() ? GT_OK : GT_ERR_PARSE;
}

}  // namespace gt

extern "C" {

int gt_wire_parse_extender_args(const char *json, size_t len, uint32_t *out_gpu_count, char *out_uid, size_t uid_cap,
                                char *out_names, size_t names_cap, int32_t *out_n_names)
{
    gt::ExtenderArgsWire a;
    int rc = gt::parse_extender_args(json, len, &a);
    if (rc != GT_OK) return rc;
    if (out_gpu_count) *out_gpu_count = a.gpu_count;
    if (out_n_names) *out_n_names = a.has_names ? (int32_t)a.names.size() : -1;
    if ((rc = copy_out(a.uid, out_uid, uid_cap)) != GT_OK) return rc;
    std::string joined;
    for (size_t i = 0; i < a.names.size(); i++) {
        if (a.names[i].find('\n') != std::string::npos) return GT_ERR_PARSE;   // not a node name
        if (i) joined.push_back('\n');
        joined += a.names[i];
    }
    return copy_out(joined, out_names, names_cap);
}

int gt_wire_host_priorities(const char *names, uint32_t n, const uint8_t *scores, char *out, size_t cap, size_t *out_len)
{
    if ((n && (!names || !scores)) || !out) return GT_ERR_INVALID;
    std::string s = "[";
    const char *p = names;
    for (uint32_t i = 0; i < n; i++) {
        const char *e = strchr(p, '\n');
        const std::string name = e ? std::string(p, (size_t)(e - p)) : std::string(p);
        if (!e && i + 1 < n) return GT_ERR_INVALID;   // fewer names than scores
        if (i) s.push_back(',');
        s += "{\"Host\":";
        gt::json_escape(name, &s);
        s += ",\"Score\":" + std::to_string((unsigned)scores[i]) + "}";
        p = e ? e + 1 : p + name.size();
    }
    s.push_back(']');
    if (out_len) *out_len = s.size();
    return copy_out(s, out, cap);
}

int gt_wire_parse_binding_args(const char *json, size_t len, char *out_pod_name, size_t name_cap, char *out_pod_namespace,
                               size_t ns_cap, char *out_pod_uid, size_t uid_cap, char *out_node, size_t node_cap)
{
    gt::BindingArgsWire b;
    int rc = gt::parse_binding_args(json, len, &b);
    if (rc != GT_OK) return rc;
    if ((rc = copy_out(b.pod_name, out_pod_name, name_cap)) != GT_OK) return rc;
    if ((rc = copy_out(b.pod_ns, out_pod_namespace, ns_cap)) != GT_OK) return rc;
    if ((rc = copy_out(b.pod_uid, out_pod_uid, uid_cap)) != GT_OK) return rc;
    return copy_out(b.node, out_node, node_cap);
}

int gt_wire_binding_result(const char *error, char *out, size_t cap)
{
    if (!out) return GT_ERR_INVALID;
    std::string s = "{\"Error\":";
    gt::json_escape(error ? error : "", &s);
    s.push_back('}');
    return copy_out(s, out, cap);
}

}  // extern "C"
