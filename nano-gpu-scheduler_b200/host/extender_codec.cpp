// extender_codec.cpp — see extender_codec.h.  A purpose-built scanner for the three request
// shapes the extender receives and a writer for the three it answers with; the rules are Go 1.16
// encoding/json's (decode.go / encode.go / scanner.go), restated, not a general JSON library.
#include "extender_codec.h"

#include <algorithm>
#include <chrono>
#include <cstring>
#include <numeric>

namespace ngs_host {

namespace {

// ---------------------------------------------------------------------------------------------
// scanner
// ---------------------------------------------------------------------------------------------
struct Scanner {
    const char* p;
    const char* end;
    Error err;      // first syntax error (decoding stops)
    Error type_err; // first type mismatch (decoding continues, reported at the end like Go's d.saveError)

    bool failed() const { return !err.empty(); }

    static std::string quote_char(unsigned char c) {  // scanner.go quoteChar
        if (c == '\'') return "'\\''";
        if (c == '"') return "'\"'";
        char buf[16];
        if (c >= 0x20 && c < 0x7f) {
            snprintf(buf, sizeof buf, "'%c'", c);
        } else if (c == '\n' || c == '\r' || c == '\t') {
            snprintf(buf, sizeof buf, "'\\%c'", c == '\n' ? 'n' : c == '\r' ? 'r' : 't');
        } else if (c < 0x20 || c == 0x7f) {
            snprintf(buf, sizeof buf, "'\\x%02x'", c);
        } else {
            snprintf(buf, sizeof buf, "'\\u00%02x'", c);  // Go quotes the byte as a rune
        }
        return buf;
    }
    void syntax(const char* context) {
        if (!err.empty()) return;
        if (p >= end) err = "unexpected EOF";
        else err = "invalid character " + quote_char((unsigned char)*p) + " " + context;
    }
    void ws() {
        while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++;
    }
    // after ws(): the next byte, or 0 at the end (with the EOF error set)
    char peek(const char* context) {
        ws();
        if (p >= end || *p == 0) {  // (a NUL byte is an invalid character wherever a token is expected)
            syntax(context);
            return 0;
        }
        return *p;
    }
    bool literal(const char* word, const char* name) {
        // scanner.go: "in literal true (expecting 'r')"
        for (size_t i = 0; word[i]; i++) {
            if (p + i >= end) {
                p = end;
                syntax("");
                return false;
            }
            if (p[i] != word[i]) {
                p += i;
                err = "invalid character " + quote_char((unsigned char)*p) + " in literal " + name + " (expecting " +
                      quote_char((unsigned char)word[i]) + ")";
                return false;
            }
        }
        p += strlen(word);
        return true;
    }
    bool number() {  // scanner.go state1/state0/dot/e...: -?(0|[1-9][0-9]*)(\.[0-9]+)?([eE][+-]?[0-9]+)?
        if (p < end && *p == '-') {
            p++;
            if (p >= end || *p < '0' || *p > '9') { syntax("in numeric literal"); return false; }
        }
        if (*p == '0') p++;
        else while (p < end && *p >= '0' && *p <= '9') p++;
        if (p < end && *p == '.') {
            p++;
            if (p >= end || *p < '0' || *p > '9') { syntax("after decimal point in numeric literal"); return false; }
            while (p < end && *p >= '0' && *p <= '9') p++;
        }
        if (p < end && (*p == 'e' || *p == 'E')) {
            p++;
            if (p < end && (*p == '+' || *p == '-')) p++;
            if (p >= end || *p < '0' || *p > '9') { syntax("in exponent of numeric literal"); return false; }
            while (p < end && *p >= '0' && *p <= '9') p++;
        }
        return true;
    }

    // byte classes of a string body: 1 = ends the fast run ('"', '\\', control), 2 = needs unescaping
    // on input or escaping on output (>= 0x80), 4 = escaped on output only ('<', '>', '&')
    static const uint8_t* classes() {
        static uint8_t t[256];
        static bool init = false;
        if (!init) {
            for (int c = 0; c < 256; c++) t[c] = c < 0x20 ? 1 : c >= 0x80 ? 2 : 0;
            t[(unsigned char)'"'] = 1;
            t[(unsigned char)'\\'] = 1;
            t[(unsigned char)'<'] = t[(unsigned char)'>'] = t[(unsigned char)'&'] = 4;
            init = true;
        }
        return t;
    }

    // p at the opening quote.  Returns the string's bytes between the quotes, whether they can be used
    // as they are (`plain`: no backslash, no byte >= 0x80) and whether json.Marshal would write the very
    // same token back (`verbatim`: plain and none of < > &); leaves p after the closing quote.
    bool string_raw(std::string_view* raw, bool* plain, bool* verbatim = nullptr) {
        static const uint8_t* T = classes();
        const char* s = ++p;
        bool pl = true;
        unsigned seen = 0;
        for (;;) {
            // find the next quote, backslash or control byte
            while (p < end) {
                const unsigned k = T[(unsigned char)*p];
                if (k & 1) break;
                seen |= k;
                p++;
            }
            if (seen & 2) pl = false;
            if (p >= end) { syntax(""); return false; }
            if (*p == '"') break;
            if ((unsigned char)*p < 0x20) { syntax("in string literal"); return false; }
            // backslash
            pl = false;
            p++;
            if (p >= end) { syntax(""); return false; }
            switch (*p) {
                case '"': case '\\': case '/': case 'b': case 'f': case 'n': case 'r': case 't':
                    p++;
                    break;
                case 'u':
                    p++;
                    for (int i = 0; i < 4; i++, p++) {
                        if (p >= end) { syntax(""); return false; }
                        const char c = *p;
                        if (!((c >= '0' && c <= '9') || (c >= 'a' && c <= 'f') || (c >= 'A' && c <= 'F'))) {
                            syntax("in \\u hexadecimal character escape");
                            return false;
                        }
                    }
                    break;
                default:
                    syntax("in string escape code");
                    return false;
            }
        }
        *raw = std::string_view(s, (size_t)(p - s));
        *plain = pl;
        if (verbatim) *verbatim = pl && !(seen & 4);
        p++;
        return true;
    }

    // skip any value (unknown keys, ignored sub-trees); *kind gets the Go name of what was there
    bool skip(const char** kind = nullptr, int depth = 0) {
        if (depth > 1000) { err = "exceeded max depth"; return false; }  // (Go allows 10000: its stacks grow, this one does not)
        const char c = peek("looking for beginning of value");
        if (!c) return false;
        const char* k = "";
        bool ok = true;
        if (c == '{') {
            k = "object";
            p++;
            if (peek("looking for beginning of object key string") == '}') { p++; }
            else for (;;) {
                if (peek("looking for beginning of object key string") != '"') { syntax("looking for beginning of object key string"); return false; }
                std::string_view r; bool pl;
                if (!string_raw(&r, &pl)) return false;
                if (peek("after object key") != ':') { syntax("after object key"); return false; }
                p++;
                if (!skip(nullptr, depth + 1)) return false;
                const char d = peek("after object key:value pair");
                if (d == ',') { p++; continue; }
                if (d == '}') { p++; break; }
                syntax("after object key:value pair");
                return false;
            }
        } else if (c == '[') {
            k = "array";
            p++;
            if (peek("looking for beginning of value") == ']') { p++; }
            else for (;;) {
                if (!skip(nullptr, depth + 1)) return false;
                const char d = peek("after array element");
                if (d == ',') { p++; continue; }
                if (d == ']') { p++; break; }
                syntax("after array element");
                return false;
            }
        } else if (c == '"') {
            k = "string";
            std::string_view r; bool pl;
            ok = string_raw(&r, &pl);
        } else if (c == 't') { k = "bool"; ok = literal("true", "true"); }
        else if (c == 'f') { k = "bool"; ok = literal("false", "false"); }
        else if (c == 'n') { k = "null"; ok = literal("null", "null"); }
        else if (c == '-' || (c >= '0' && c <= '9')) { k = "number"; ok = number(); }
        else { syntax("looking for beginning of value"); return false; }
        if (kind) *kind = k;
        return ok;
    }
};

void append_utf8(std::string* out, uint32_t cp) {
    if (cp < 0x80) out->push_back((char)cp);
    else if (cp < 0x800) { out->push_back((char)(0xC0 | (cp >> 6))); out->push_back((char)(0x80 | (cp & 0x3F))); }
    else if (cp < 0x10000) {
        out->push_back((char)(0xE0 | (cp >> 12))); out->push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out->push_back((char)(0x80 | (cp & 0x3F)));
    } else {
        out->push_back((char)(0xF0 | (cp >> 18))); out->push_back((char)(0x80 | ((cp >> 12) & 0x3F)));
        out->push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out->push_back((char)(0x80 | (cp & 0x3F)));
    }
}

// length of the valid UTF-8 sequence at s (0 = invalid), as utf8.DecodeRune accepts it
int utf8_len(const unsigned char* s, size_t n, uint32_t* cp) {
    const unsigned char c = s[0];
    if (c < 0x80) { *cp = c; return 1; }
    if (c < 0xC2) return 0;
    if (c < 0xE0) {
        if (n < 2 || (s[1] & 0xC0) != 0x80) return 0;
        *cp = ((c & 0x1F) << 6) | (s[1] & 0x3F);
        return 2;
    }
    if (c < 0xF0) {
        if (n < 3 || (s[1] & 0xC0) != 0x80 || (s[2] & 0xC0) != 0x80) return 0;
        if (c == 0xE0 && s[1] < 0xA0) return 0;
        if (c == 0xED && s[1] > 0x9F) return 0;  // surrogates
        *cp = ((c & 0x0F) << 12) | ((s[1] & 0x3F) << 6) | (s[2] & 0x3F);
        return 3;
    }
    if (c < 0xF5) {
        if (n < 4 || (s[1] & 0xC0) != 0x80 || (s[2] & 0xC0) != 0x80 || (s[3] & 0xC0) != 0x80) return 0;
        if (c == 0xF0 && s[1] < 0x90) return 0;
        if (c == 0xF4 && s[1] > 0x8F) return 0;
        *cp = ((c & 0x07) << 18) | ((s[1] & 0x3F) << 12) | ((s[2] & 0x3F) << 6) | (s[3] & 0x3F);
        return 4;
    }
    return 0;
}

int hex4(const char* p) {
    int v = 0;
    for (int i = 0; i < 4; i++) {
        const char c = p[i];
        v = v * 16 + (c <= '9' ? c - '0' : (c | 0x20) - 'a' + 10);
    }
    return v;
}

// decode.go unquoteBytes: escapes resolved, lone surrogates and invalid UTF-8 become U+FFFD
std::string unquote(std::string_view raw) {
    std::string out;
    out.reserve(raw.size());
    const char* p = raw.data();
    const char* end = p + raw.size();
    while (p < end) {
        const unsigned char c = (unsigned char)*p;
        if (c == '\\') {
            p++;
            switch (*p) {
                case '"': out.push_back('"'); p++; break;
                case '\\': out.push_back('\\'); p++; break;
                case '/': out.push_back('/'); p++; break;
                case 'b': out.push_back('\b'); p++; break;
                case 'f': out.push_back('\f'); p++; break;
                case 'n': out.push_back('\n'); p++; break;
                case 'r': out.push_back('\r'); p++; break;
                case 't': out.push_back('\t'); p++; break;
                default: {  // 'u' (string_raw validated the shape)
                    uint32_t cp = (uint32_t)hex4(p + 1);
                    p += 5;
                    if (cp >= 0xD800 && cp < 0xDC00) {
                        if (end - p >= 6 && p[0] == '\\' && p[1] == 'u') {
                            const uint32_t lo = (uint32_t)hex4(p + 2);
                            if (lo >= 0xDC00 && lo < 0xE000) {
                                cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                                p += 6;
                            } else cp = 0xFFFD;
                        } else cp = 0xFFFD;
                    } else if (cp >= 0xDC00 && cp < 0xE000) cp = 0xFFFD;
                    append_utf8(&out, cp);
                }
            }
        } else if (c < 0x80) {
            out.push_back((char)c);
            p++;
        } else {
            uint32_t cp;
            const int n = utf8_len((const unsigned char*)p, (size_t)(end - p), &cp);
            if (n == 0) {
                append_utf8(&out, 0xFFFD);
                p++;
            } else {
                out.append(p, (size_t)n);
                p += n;
            }
        }
    }
    return out;
}

bool key_is(std::string_view raw, bool plain, const char* name) {  // decode.go: exact, else case-folded
    if (plain) {
        const size_t n = strlen(name);
        if (raw.size() != n) return false;
        for (size_t i = 0; i < n; i++) {
            char a = raw[i], b = name[i];
            if (a >= 'A' && a <= 'Z') a |= 0x20;
            if (b >= 'A' && b <= 'Z') b |= 0x20;
            if (a != b) return false;
        }
        return true;
    }
    const std::string u = unquote(raw);
    return key_is(u, true, name);
}

void type_mismatch(Scanner& s, const char* got, const char* field, const char* type) {
    if (s.type_err.empty() && strcmp(got, "null") != 0)
        s.type_err = std::string("json: cannot unmarshal ") + got + " into Go struct field " + field + " of type " + type;
}

// Iterate the members of an object: calls f(key_raw, key_plain) with p at the value; f must consume it.
template <class F>
bool object_members(Scanner& s, F f) {
    s.p++;  // '{'
    if (s.peek("looking for beginning of object key string") == '}') { s.p++; return true; }
    for (;;) {
        if (s.peek("looking for beginning of object key string") != '"') { s.syntax("looking for beginning of object key string"); return false; }
        std::string_view k; bool pl;
        if (!s.string_raw(&k, &pl)) return false;
        if (s.peek("after object key") != ':') { s.syntax("after object key"); return false; }
        s.p++;
        if (!s.peek("looking for beginning of value")) return false;
        if (!f(k, pl)) return false;
        const char d = s.peek("after object key:value pair");
        if (d == ',') { s.p++; continue; }
        if (d == '}') { s.p++; return true; }
        s.syntax("after object key:value pair");
        return false;
    }
}

// a value expected to be an object: runs `f` over its members; null is a no-op; anything else is a type mismatch
template <class F>
bool expect_object(Scanner& s, const char* field, const char* type, F f) {
    const char c = s.peek("looking for beginning of value");
    if (!c) return false;
    if (c == '{') return object_members(s, f);
    const char* kind;
    if (!s.skip(&kind)) return false;
    type_mismatch(s, kind, field, type);
    return true;
}

// a value expected to be a string; *set says whether one was stored (null leaves the target alone)
bool expect_string(Scanner& s, const char* field, std::string* out, bool* set = nullptr) {
    const char c = s.peek("looking for beginning of value");
    if (!c) return false;
    if (set) *set = false;
    if (c == '"') {
        std::string_view raw; bool pl;
        if (!s.string_raw(&raw, &pl)) return false;
        *out = pl ? std::string(raw) : unquote(raw);
        if (set) *set = true;
        return true;
    }
    const char* kind;
    if (!s.skip(&kind)) return false;
    type_mismatch(s, kind, field, "string");
    return true;
}

bool string_map(Scanner& s, const char* field, std::map<std::string, std::string>* out) {
    return expect_object(s, field, "map[string]string", [&](std::string_view k, bool pl) {
        std::string v; bool set;
        if (!expect_string(s, field, &v, &set)) return false;
        (*out)[pl ? std::string(k) : unquote(k)] = set ? v : std::string();
        return true;
    });
}

const char* kResourceGPUPercent = "nano-gpu/gpu-percent";  // pkg/types/types.go:9

bool decode_limits(Scanner& s, Container* c) {
    return expect_object(s, "ResourceRequirements.spec.containers.resources.limits", "v1.ResourceList", [&](std::string_view k, bool pl) {
        const bool mine = pl ? k == kResourceGPUPercent : unquote(k) == kResourceGPUPercent;  // a Go map key: exact match
        if (!mine) return s.skip();
        // resource.Quantity.UnmarshalJSON: null -> zero; a quoted string or a bare token, trimmed, then ParseQuantity
        const char ch = *s.p;
        std::string text;
        if (ch == '"') {
            std::string_view raw; bool p2;
            if (!s.string_raw(&raw, &p2)) return false;
            text = p2 ? std::string(raw) : unquote(raw);
        } else {
            const char* b = s.p;
            const char* kind;
            if (!s.skip(&kind)) return false;
            text.assign(b, (size_t)(s.p - b));
            if (text == "null") {
                c->has_gpu_limit = true;
                c->gpu_percent = 0;
                return true;
            }
        }
        int64_t v;
        if (!ParseQuantityValue(text, &v)) {
            if (s.type_err.empty())
                s.type_err = "quantities must match the regular expression '^([+-]?[0-9.]+)([eEinumkKMGTP]*[-+]?[0-9]*)$'";
            return true;
        }
        c->has_gpu_limit = true;
        c->gpu_percent = v;
        return true;
    });
}

bool decode_container(Scanner& s, Container* c) {
    return expect_object(s, "PodSpec.spec.containers", "v1.Container", [&](std::string_view k, bool pl) {
        if (key_is(k, pl, "name")) return expect_string(s, "Container.spec.containers.name", &c->name);
        if (key_is(k, pl, "resources"))
            return expect_object(s, "Container.spec.containers.resources", "v1.ResourceRequirements", [&](std::string_view k2, bool pl2) {
                if (key_is(k2, pl2, "limits")) return decode_limits(s, c);
                return s.skip();
            });
        return s.skip();
    });
}

bool decode_pod(Scanner& s, Pod* pod) {
    return expect_object(s, "ExtenderArgs.Pod", "v1.Pod", [&](std::string_view k, bool pl) {
        if (key_is(k, pl, "metadata"))
            return expect_object(s, "Pod.metadata", "v1.ObjectMeta", [&](std::string_view k2, bool pl2) {
                if (key_is(k2, pl2, "name")) return expect_string(s, "ObjectMeta.metadata.name", &pod->name);
                if (key_is(k2, pl2, "namespace")) return expect_string(s, "ObjectMeta.metadata.namespace", &pod->ns);
                if (key_is(k2, pl2, "uid")) return expect_string(s, "ObjectMeta.metadata.uid", &pod->uid);
                if (key_is(k2, pl2, "annotations")) return string_map(s, "ObjectMeta.metadata.annotations", &pod->annotations);
                if (key_is(k2, pl2, "labels")) return string_map(s, "ObjectMeta.metadata.labels", &pod->labels);
                return s.skip();
            });
        if (key_is(k, pl, "spec"))
            return expect_object(s, "Pod.spec", "v1.PodSpec", [&](std::string_view k2, bool pl2) {
                if (key_is(k2, pl2, "nodeName")) return expect_string(s, "PodSpec.spec.nodeName", &pod->node_name);
                if (key_is(k2, pl2, "containers")) {
                    const char c = *s.p;
                    if (c != '[') {
                        const char* kind;
                        if (!s.skip(&kind)) return false;
                        type_mismatch(s, kind, "PodSpec.spec.containers", "[]v1.Container");
                        if (strcmp(kind, "null") == 0) pod->containers.clear();
                        return true;
                    }
                    pod->containers.clear();
                    s.p++;
                    if (s.peek("looking for beginning of value") == ']') { s.p++; return true; }
                    for (;;) {
                        pod->containers.emplace_back();
                        if (!decode_container(s, &pod->containers.back())) return false;
                        const char d = s.peek("after array element");
                        if (d == ',') { s.p++; continue; }
                        if (d == ']') { s.p++; return true; }
                        s.syntax("after array element");
                        return false;
                    }
                }
                return s.skip();
            });
        return s.skip();
    });
}

bool decode_node_names(Scanner& s, ExtenderArgs* out, const std::string* cached) {
    const char c = *s.p;
    if (cached && !cached->empty() && (size_t)(s.end - s.p) >= cached->size() && memcmp(s.p, cached->data(), cached->size()) == 0) {
        // byte for byte the array that was tokenised (and found well-formed) last time: a JSON array is
        // self-delimiting, so the same bytes from the same state parse to the same list
        out->has_node_names = true;
        out->node_names_cached = true;
        out->node_names.clear();
        out->name_verbatim.clear();
        out->node_names_raw = std::string_view(s.p, cached->size());
        s.p += cached->size();
        return true;
    }
    out->node_names_cached = false;
    if (c != '[') {
        const char* kind;
        if (!s.skip(&kind)) return false;
        if (strcmp(kind, "null") == 0) {  // a nil *[]string
            out->has_node_names = false;
            out->node_names.clear();
            out->name_verbatim.clear();
        } else type_mismatch(s, kind, "ExtenderArgs.NodeNames", "[]string");
        return true;
    }
    const char* begin = s.p;
    out->has_node_names = true;
    out->node_names.clear();
    out->name_verbatim.clear();
    s.p++;
    if (s.peek("looking for beginning of value") == ']') {
        s.p++;
    } else {
        for (;;) {
            const char e = s.peek("looking for beginning of value");
            if (!e) return false;
            if (e == '"') {
                std::string_view raw; bool pl, vb;
                if (!s.string_raw(&raw, &pl, &vb)) return false;
                if (pl) out->node_names.push_back(raw);
                else {
                    out->unescaped.push_back(unquote(raw));
                    out->node_names.push_back(out->unescaped.back());
                }
                out->name_verbatim.push_back(vb ? 1 : 0);
            } else {
                const char* kind;
                if (!s.skip(&kind)) return false;
                type_mismatch(s, kind, "ExtenderArgs.NodeNames", "string");
                out->node_names.push_back(std::string_view());  // the element keeps its zero value
                out->name_verbatim.push_back(0);
            }
            const char d = s.peek("after array element");
            if (d == ',') { s.p++; continue; }
            if (d == ']') { s.p++; break; }
            s.syntax("after array element");
            return false;
        }
    }
    out->node_names_raw = std::string_view(begin, (size_t)(s.p - begin));
    return true;
}

Error top_level(std::string_view body, Scanner* s) {
    s->p = body.data();
    s->end = body.data() + body.size();
    s->ws();
    if (s->p >= s->end) return "EOF";  // json.Decoder on an empty stream
    return "";
}

}  // namespace

Error DecodeExtenderArgs(std::string_view body, ExtenderArgs* out, const std::string* cached_node_names) {
    Scanner s;
    Error e = top_level(body, &s);
    if (!e.empty()) return e;
    const char c = *s.p;
    if (c == '{') {
        object_members(s, [&](std::string_view k, bool pl) {
            if (key_is(k, pl, "Pod")) {
                if (*s.p == '{') out->has_pod = true;
                else if (*s.p == 'n') out->has_pod = false, out->pod = Pod();
                return decode_pod(s, &out->pod);
            }
            if (key_is(k, pl, "NodeNames")) return decode_node_names(s, out, cached_node_names);
            if (key_is(k, pl, "Nodes")) {
                const char* kind;
                if (!s.skip(&kind)) return false;
                out->has_nodes = strcmp(kind, "null") != 0;
                if (strcmp(kind, "object") != 0) type_mismatch(s, kind, "ExtenderArgs.Nodes", "v1.NodeList");
                return true;
            }
            return s.skip();
        });
    } else {
        const char* kind;
        if (s.skip(&kind) && strcmp(kind, "null") != 0 && s.type_err.empty())
            s.type_err = std::string("json: cannot unmarshal ") + kind + " into Go value of type v1.ExtenderArgs";
    }
    if (s.failed()) return s.err;
    return s.type_err;
}

Error DecodeBindingArgs(std::string_view body, BindingArgs* out) {
    Scanner s;
    Error e = top_level(body, &s);
    if (!e.empty()) return e;
    if (*s.p == '{') {
        object_members(s, [&](std::string_view k, bool pl) {
            if (key_is(k, pl, "PodName")) return expect_string(s, "ExtenderBindingArgs.PodName", &out->PodName);
            if (key_is(k, pl, "PodNamespace")) return expect_string(s, "ExtenderBindingArgs.PodNamespace", &out->PodNamespace);
            if (key_is(k, pl, "PodUID")) return expect_string(s, "ExtenderBindingArgs.PodUID", &out->PodUID);
            if (key_is(k, pl, "Node")) return expect_string(s, "ExtenderBindingArgs.Node", &out->Node);
            return s.skip();
        });
    } else {
        const char* kind;
        if (s.skip(&kind) && strcmp(kind, "null") != 0 && s.type_err.empty())
            s.type_err = std::string("json: cannot unmarshal ") + kind + " into Go value of type v1.ExtenderBindingArgs";
    }
    if (s.failed()) return s.err;
    return s.type_err;
}

// ---------------------------------------------------------------------------------------------
// resource.Quantity
// ---------------------------------------------------------------------------------------------
bool ParseQuantityValue(std::string_view s, int64_t* out) {
    // apimachinery/pkg/api/resource/quantity.go parseQuantityString + Value(): <sign><digits>[.<digits>]<suffix>,
    // suffix one of Ki Mi Gi Ti Pi Ei | n u m "" k M G T P E | e<int> E<int>; the value is rounded up.
    while (!s.empty() && (s.front() == ' ' || s.front() == '\t' || s.front() == '\n' || s.front() == '\r')) s.remove_prefix(1);
    while (!s.empty() && (s.back() == ' ' || s.back() == '\t' || s.back() == '\n' || s.back() == '\r')) s.remove_suffix(1);
    if (s.empty()) return false;
    size_t i = 0;
    bool neg = false;
    if (s[i] == '+' || s[i] == '-') neg = s[i++] == '-';
    unsigned __int128 mant = 0;
    int exp10 = 0, digits = 0;
    bool overflow = false;
    auto push = [&](char c) {
        if (mant > ((unsigned __int128)1 << 100)) overflow = true;
        else mant = mant * 10 + (unsigned)(c - '0');
        digits++;
    };
    while (i < s.size() && s[i] >= '0' && s[i] <= '9') push(s[i++]);
    if (i < s.size() && s[i] == '.') {
        i++;
        while (i < s.size() && s[i] >= '0' && s[i] <= '9') {
            push(s[i++]);
            exp10--;
        }
    }
    if (digits == 0) return false;
    const std::string_view suf = s.substr(i);
    int exp2 = 0;
    if (suf.empty()) {
    } else if (suf == "Ki") exp2 = 10;
    else if (suf == "Mi") exp2 = 20;
    else if (suf == "Gi") exp2 = 30;
    else if (suf == "Ti") exp2 = 40;
    else if (suf == "Pi") exp2 = 50;
    else if (suf == "Ei") exp2 = 60;
    else if (suf == "n") exp10 -= 9;
    else if (suf == "u") exp10 -= 6;
    else if (suf == "m") exp10 -= 3;
    else if (suf == "k") exp10 += 3;
    else if (suf == "M") exp10 += 6;
    else if (suf == "G") exp10 += 9;
    else if (suf == "T") exp10 += 12;
    else if (suf == "P") exp10 += 15;
    else if (suf == "E") exp10 += 18;
    else if (suf[0] == 'e' || suf[0] == 'E') {
        size_t j = 1;
        bool eneg = false;
        if (j < suf.size() && (suf[j] == '+' || suf[j] == '-')) eneg = suf[j++] == '-';
        if (j >= suf.size()) return false;
        int ev = 0;
        for (; j < suf.size(); j++) {
            if (suf[j] < '0' || suf[j] > '9') return false;
            if (ev < 100000) ev = ev * 10 + (suf[j] - '0');
        }
        exp10 += eneg ? -ev : ev;
    } else return false;

    // value = mant * 2^exp2 * 10^exp10, rounded up, saturated at int64.  With k = -exp10 > 0 it is split as
    // mant = q * 10^k + r, value = q * 2^exp2 + ceil(r * 2^exp2 / 10^k), the last term with the common powers of
    // two cancelled so that everything stays inside 128 bits (k is held to 28 digits; more fractional digits
    // than that only keep a "something was there" flag, which is all a ceiling needs for them).
    const unsigned __int128 kMax = (unsigned __int128)INT64_MAX;
    if (mant == 0 && !overflow) {
        *out = 0;
        return true;
    }
    unsigned __int128 v;
    bool sticky = false;
    if (overflow) {
        v = kMax + 1;
    } else if (exp10 >= 0) {
        v = mant;
        for (int k = 0; k < exp2 && v <= kMax; k++) v *= 2;
        for (; exp10 > 0 && v <= kMax; exp10--) v *= 10;
    } else {
        int k = -exp10;
        for (; k > 28; k--) {
            if (mant % 10) sticky = true;
            mant /= 10;
        }
        unsigned __int128 D = 1;
        for (int t = 0; t < k; t++) D *= 10;
        const unsigned __int128 q = mant / D, r = mant % D;
        v = q;
        for (int t = 0; t < exp2 && v <= kMax; t++) v *= 2;
        const int c = exp2 < k ? exp2 : k;  // twos shared by 2^exp2 and 10^k
        unsigned __int128 Dr = D, num = r;
        for (int t = 0; t < c; t++) Dr /= 2;
        for (int t = 0; t < exp2 - c; t++) num *= 2;  // r < 10^k, so num < 5^k * 2^exp2 <= 5^28 * 2^60 < 2^126
        unsigned __int128 frac = num / Dr;
        if (num % Dr) sticky = true;
        if (v <= kMax) v += frac;
    }
    if (sticky && !neg && v <= kMax) v += 1;  // Value() = ScaledValue(0) rounds up
    if (v > kMax) v = kMax;                   // Value() saturates
    *out = neg ? -(int64_t)v : (int64_t)v;
    return true;
}

// ---------------------------------------------------------------------------------------------
// encoder
// ---------------------------------------------------------------------------------------------
void AppendJSONString(std::string* out, std::string_view s) {  // encode.go (*encodeState).string, escapeHTML = true
    static const char* hex = "0123456789abcdef";
    out->push_back('"');
    const unsigned char* p = (const unsigned char*)s.data();
    const size_t n = s.size();
    size_t start = 0, i = 0;
    while (i < n) {
        const unsigned char c = p[i];
        if (c < 0x80) {
            if (c >= 0x20 && c != '"' && c != '\\' && c != '<' && c != '>' && c != '&') {
                i++;
                continue;
            }
            out->append((const char*)p + start, i - start);
            switch (c) {
                case '"': out->append("\\\""); break;
                case '\\': out->append("\\\\"); break;
                case '\n': out->append("\\n"); break;
                case '\r': out->append("\\r"); break;
                case '\t': out->append("\\t"); break;
                default:
                    out->append("\\u00");
                    out->push_back(hex[c >> 4]);
                    out->push_back(hex[c & 0xF]);
            }
            start = ++i;
            continue;
        }
        uint32_t cp;
        const int len = utf8_len(p + i, n - i, &cp);
        if (len == 0) {
            out->append((const char*)p + start, i - start);
            out->append("\\ufffd");
            start = ++i;
            continue;
        }
        if (cp == 0x2028 || cp == 0x2029) {
            out->append((const char*)p + start, i - start);
            out->append(cp == 0x2028 ? "\\u2028" : "\\u2029");
            i += (size_t)len;
            start = i;
            continue;
        }
        i += (size_t)len;
    }
    out->append((const char*)p + start, n - start);
    out->push_back('"');
}

std::string EncodeFilterResult(const std::vector<std::string_view>& can, const std::vector<std::pair<std::string_view, std::string>>& failed) {
    size_t bytes = 64;
    for (auto& s : can) bytes += s.size() + 3;
    for (auto& kv : failed) bytes += kv.first.size() + kv.second.size() + 6;
    std::string out;
    out.reserve(bytes);
    out += "{\"Nodes\":null,\"NodeNames\":[";
    for (size_t i = 0; i < can.size(); i++) {
        if (i) out.push_back(',');
        AppendJSONString(&out, can[i]);
    }
    out += "],\"FailedNodes\":{";
    for (size_t i = 0; i < failed.size(); i++) {
        if (i) out.push_back(',');
        AppendJSONString(&out, failed[i].first);
        out.push_back(':');
        AppendJSONString(&out, failed[i].second);
    }
    out += "},\"Error\":\"\"}";
    return out;
}

std::string EncodeFilterError(const std::string& error) {
    std::string out = "{\"Nodes\":null,\"NodeNames\":null,\"FailedNodes\":null,\"Error\":";
    AppendJSONString(&out, error);
    out.push_back('}');
    return out;
}

std::string EncodeBindingResult(const std::string& error) {
    std::string out = "{\"Error\":";
    AppendJSONString(&out, error);
    out.push_back('}');
    return out;
}

std::string EncodeHostPriorityList(const std::vector<std::string_view>& hosts, const std::vector<int>& scores) {
    size_t bytes = 8;
    for (auto& s : hosts) bytes += s.size() + 32;
    std::string out;
    out.reserve(bytes);
    out.push_back('[');
    char num[24];
    for (size_t i = 0; i < hosts.size(); i++) {
        if (i) out.push_back(',');
        out += "{\"Host\":";
        AppendJSONString(&out, hosts[i]);
        out += ",\"Score\":";
        snprintf(num, sizeof num, "%d", scores[i]);
        out += num;
        out.push_back('}');
    }
    out.push_back(']');
    return out;
}

// ---------------------------------------------------------------------------------------------
// routes
// ---------------------------------------------------------------------------------------------
namespace {
uint64_t now_ns() { return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
}  // namespace

bool ExtenderRoutes::prepare(const ExtenderArgs& a) {
    c_.requests++;
    if (a.node_names_cached) {
        c_.cache_hits++;
        if (set_.n_missing) d_->Resolve(&set_);  // the reference asks the lister again on every request (dealer.go:100)
        return true;
    }
    valid_ = false;
    // keep the array's bytes; names and output tokens of ordinary names are views into that copy
    const char* body_raw = a.node_names_raw.data();
    raw_.assign(body_raw, a.node_names_raw.size());
    const size_t n = a.node_names.size();
    names_.resize(n);
    tok_.resize(n);
    names_store_.clear();
    tok_store_.clear();
    std::vector<std::pair<size_t, size_t>> odd;  // (name offset, token offset) of the names that are not verbatim
    tok_bytes_ = 0;
    for (size_t i = 0; i < n; i++) {
        const std::string_view nm = a.node_names[i];
        if (a.name_verbatim[i]) {
            const size_t off = (size_t)(nm.data() - body_raw);
            names_[i] = std::string_view(raw_.data() + off, nm.size());
            tok_[i] = std::string_view(raw_.data() + off - 1, nm.size() + 2);
        } else {
            odd.emplace_back(names_store_.size(), tok_store_.size());
            names_store_.append(nm.data(), nm.size());
            AppendJSONString(&tok_store_, nm);
            names_[i] = nm;  // re-pointed below, once the stores stop moving
        }
    }
    size_t k = 0;
    for (size_t i = 0; i < n; i++) {
        if (!a.name_verbatim[i]) {
            const size_t tok_end = k + 1 < odd.size() ? odd[k + 1].second : tok_store_.size();
            names_[i] = std::string_view(names_store_.data() + odd[k].first, names_[i].size());
            tok_[i] = std::string_view(tok_store_.data() + odd[k].second, tok_end - odd[k].second);
            k++;
        }
        tok_bytes_ += tok_[i].size();
    }
    set_ = CandidateSet();
    set_.names = names_;
    d_->Resolve(&set_);
    order_.clear();  // built by the first filter answer that needs it
    valid_ = true;
    return false;
}

namespace {
inline char* put(char* p, std::string_view s) {
    memcpy(p, s.data(), s.size());
    return p + s.size();
}
inline char* put_int(char* p, int64_t v) {
    char tmp[24];
    int n = 0;
    uint64_t u = v < 0 ? (uint64_t)0 - (uint64_t)v : (uint64_t)v;
    do tmp[n++] = (char)('0' + u % 10); while (u /= 10);
    if (v < 0) *p++ = '-';
    while (n) *p++ = tmp[--n];
    return p;
}
}  // namespace

std::string ExtenderRoutes::Predicate(std::string_view body) {
    std::lock_guard<std::mutex> lock(mu_);
    const uint64_t t0 = now_ns();
    ExtenderArgs a;
    Error e = DecodeExtenderArgs(body, &a, valid_ ? &raw_ : nullptr);
    const uint64_t t1 = now_ns();
    c_.ns_decode += t1 - t0;
    if (!e.empty()) return EncodeFilterError(e);  // routes.go:54-60
    if (!a.has_node_names) return EncodeFilterError("nano-gpu-scheduler extender must be configured with nodeCacheCapable=true");
    prepare(a);
    if (set_.all_rows) c_.all_rows++;
    const uint64_t t2 = now_ns();
    c_.ns_resolve += t2 - t1;
    AssumeResult r = d_->AssumeSet(set_, a.pod, spec_, load_);
    const uint64_t t3 = now_ns();
    c_.ns_dealer += t3 - t2;

    // json.Marshal(&ExtenderFilterResult{...}) written straight into one buffer: names go out as their
    // cached tokens, the rater's failure texts are assembled in place (plain ASCII, nothing to escape)
    const size_t n = names_.size();
    size_t n_fail = 0, slow_bytes = 0;
    for (size_t i = 0; i < n; i++) n_fail += r.ok[i] ? 0 : 1;
    std::vector<std::pair<uint32_t, std::string>> slow;  // failures whose text needs the general encoder
    if (n_fail) {
        if (order_.size() != n) {  // ascending-name order of the candidates (Go sorts map keys), kept with the cached set
            order_.resize(n);
            std::iota(order_.begin(), order_.end(), 0u);
            std::stable_sort(order_.begin(), order_.end(), [&](uint32_t x, uint32_t y) { return names_[x] < names_[y]; });
        }
        for (size_t i = 0; i < n; i++)
            if (!r.ok[i] && !d_->FailureIsPlain(set_, r, i)) {
                std::string m;
                AppendJSONString(&m, d_->FailureText(set_, r, i));
                slow_bytes += m.size();
                slow.emplace_back((uint32_t)i, std::move(m));
            }
    }
    const size_t per_msg = 64 + r.demand_text.size() + NGS_MAX_GPUS_WIDE * 8 + NGS_MAX_CONTAINERS * 12;
    std::string out;
    out.resize(96 + tok_bytes_ + n + n_fail * (per_msg + 2) + slow_bytes);
    char* p = &out[0];
    p = put(p, "{\"Nodes\":null,\"NodeNames\":[");
    bool first = true;
    for (size_t i = 0; i < n; i++) {
        if (!r.ok[i]) continue;
        if (!first) *p++ = ',';
        first = false;
        p = put(p, tok_[i]);
    }
    p = put(p, "],\"FailedNodes\":{");
    first = true;
    std::vector<uint32_t> slow_at;  // candidate index -> position in `slow`
    if (!slow.empty()) {
        slow_at.assign(n, 0xFFFFFFFFu);
        for (size_t k = 0; k < slow.size(); k++) slow_at[slow[k].first] = (uint32_t)k;
    }
    for (size_t k = 0; k < n && n_fail; k++) {
        const uint32_t i = order_[k];
        if (r.ok[i]) continue;
        // a name listed twice is one map key; the later entry's message wins (predicate.go:33)
        if (k + 1 < n && names_[order_[k + 1]] == names_[i] && !r.ok[order_[k + 1]]) continue;
        if (!first) *p++ = ',';
        first = false;
        p = put(p, tok_[i]);
        *p++ = ':';
        if (!slow_at.empty() && slow_at[i] != 0xFFFFFFFFu) p = put(p, slow[slow_at[i]].second);
        else {
            *p++ = '"';
            p = d_->PutFailureText(p, set_, r, i);
            *p++ = '"';
        }
    }
    p = put(p, "},\"Error\":\"\"}");
    out.resize((size_t)(p - out.data()));
    c_.ns_encode += now_ns() - t3;
    return out;
}

std::string ExtenderRoutes::Prioritize(std::string_view body, bool* panicked) {
    std::lock_guard<std::mutex> lock(mu_);
    if (panicked) *panicked = false;
    const uint64_t t0 = now_ns();
    ExtenderArgs a;
    Error e = DecodeExtenderArgs(body, &a, valid_ ? &raw_ : nullptr);
    const uint64_t t1 = now_ns();
    c_.ns_decode += t1 - t0;
    if (!e.empty() || !a.has_node_names) {  // routes.go:103 panic(err); priority.go:24 dereferences a nil NodeNames
        if (panicked) *panicked = true;
        return std::string();
    }
    prepare(a);
    if (set_.all_rows) c_.all_rows++;
    const uint64_t t2 = now_ns();
    c_.ns_resolve += t2 - t1;
    std::vector<int> scores = d_->ScoreSet(set_, a.pod, spec_, load_);
    const uint64_t t3 = now_ns();
    c_.ns_dealer += t3 - t2;
    const size_t n = names_.size();
    std::string out;
    out.resize(8 + tok_bytes_ + n * 40);
    char* p = &out[0];
    *p++ = '[';
    for (size_t i = 0; i < n; i++) {
        if (i) *p++ = ',';
        p = put(p, "{\"Host\":");
        p = put(p, tok_[i]);
        p = put(p, ",\"Score\":");
        p = put_int(p, scores[i]);
        *p++ = '}';
    }
    *p++ = ']';
    out.resize((size_t)(p - out.data()));
    c_.ns_encode += now_ns() - t3;
    return out;
}

std::string ExtenderRoutes::Bind(std::string_view body, int* http_status) {
    std::lock_guard<std::mutex> lock(mu_);
    BindingArgs a;
    Error err = DecodeBindingArgs(body, &a);  // routes.go:136-140
    if (err.empty()) {
        Pod pod;
        ClusterSource* src = d_->source();
        // getPod (scheduler/bind.go:61-84): a second Get when the uid differs, then give up
        bool found = src && src->GetPod(a.PodNamespace, a.PodName, &pod);
        if (found && pod.uid != a.PodUID) found = src->GetPod(a.PodNamespace, a.PodName, &pod);
        char buf[1024];
        if (!found) {
            snprintf(buf, sizeof buf, "pods \"%s\" not found", a.PodName.c_str());
            err = buf;
        } else if (pod.uid != a.PodUID) {
            snprintf(buf, sizeof buf, "pod %s in ns %s's uid is %s, and it's not equal with expected %s", a.PodName.c_str(), a.PodNamespace.c_str(),
                     pod.uid.c_str(), a.PodUID.c_str());
            err = buf;
        } else if (pod.completed) {  // bind.go:46-50
            snprintf(buf, sizeof buf, "pod %s/%s already deleted or completed", a.PodName.c_str(), a.PodNamespace.c_str());
            err = buf;
        } else {
            err = d_->Bind(a.Node, &pod, spec_, load_);  // bind.go:52
        }
    }
    if (http_status) *http_status = err.empty() ? 200 : 500;  // routes.go:147-166
    return EncodeBindingResult(err);
}

}  // namespace ngs_host
