#include "json.h"

#include <cctype>
#include <cstdio>
#include <cstdlib>

namespace dissem {
namespace host {

static bool ieq(const std::string &a, const std::string &b) {
    if (a.size() != b.size()) return false;
    for (size_t i = 0; i < a.size(); ++i)
        if (std::tolower((unsigned char)a[i]) != std::tolower((unsigned char)b[i])) return false;
    return true;
}

const Json *Json::get(const std::string &key) const {
    if (kind != Object) return nullptr;
    for (auto &kv : obj)
        if (kv.first == key) return &kv.second;
    for (auto &kv : obj)
        if (ieq(kv.first, key)) return &kv.second;
    return nullptr;
}

static void dump_str(const std::string &s, std::string &out) {
    out.push_back('"');
    for (unsigned char c : s) {
        switch (c) {
            case '"': out += "\\\""; break;
            case '\\': out += "\\\\"; break;
            case '\n': out += "\\n"; break;
            case '\r': out += "\\r"; break;
            case '\t': out += "\\t"; break;
            default:
                if (c < 0x20) {
                    char buf[8];
                    snprintf(buf, sizeof buf, "\\u%04x", c);
                    out += buf;
                } else out.push_back((char)c);
        }
    }
    out.push_back('"');
}

static void dump_into(const Json &j, std::string &out) {
    switch (j.kind) {
        case Json::Null: out += "null"; break;
        case Json::Bool: out += j.b ? "true" : "false"; break;
        case Json::Int: out += std::to_string(j.i); break;
        case Json::Double: {
            char buf[64];
            snprintf(buf, sizeof buf, "%.17g", j.d);
            out += buf;
            break;
        }
        case Json::String: dump_str(j.s, out); break;
        case Json::Array:
            out.push_back('[');
            for (size_t k = 0; k < j.arr.size(); ++k) {
                if (k) out.push_back(',');
                dump_into(j.arr[k], out);
            }
            out.push_back(']');
            break;
        case Json::Object:
            out.push_back('{');
            for (size_t k = 0; k < j.obj.size(); ++k) {
                if (k) out.push_back(',');
                dump_str(j.obj[k].first, out);
                out.push_back(':');
                dump_into(j.obj[k].second, out);
            }
            out.push_back('}');
            break;
    }
}

std::string Json::dump() const {
    std::string out;
    dump_into(*this, out);
    return out;
}

static void skip_ws(const std::string &t, size_t &p) {
    while (p < t.size() && std::isspace((unsigned char)t[p])) ++p;
}

static std::string parse_string(const std::string &t, size_t &p) {
    if (t[p] != '"') throw JsonError("expected string at offset " + std::to_string(p));
    ++p;
    std::string out;
    while (p < t.size() && t[p] != '"') {
        char c = t[p++];
        if (c == '\\') {
            if (p >= t.size()) break;
            char e = t[p++];
            switch (e) {
                case 'n': out.push_back('\n'); break;
                case 't': out.push_back('\t'); break;
                case 'r': out.push_back('\r'); break;
                case 'b': out.push_back('\b'); break;
                case 'f': out.push_back('\f'); break;
                case 'u': {
                    if (p + 4 > t.size()) throw JsonError("bad \\u escape");
                    unsigned v = (unsigned)strtoul(t.substr(p, 4).c_str(), nullptr, 16);
                    p += 4;
                    if (v < 0x80) out.push_back((char)v);
                    else if (v < 0x800) { out.push_back((char)(0xC0 | (v >> 6))); out.push_back((char)(0x80 | (v & 0x3F))); }
                    else { out.push_back((char)(0xE0 | (v >> 12))); out.push_back((char)(0x80 | ((v >> 6) & 0x3F))); out.push_back((char)(0x80 | (v & 0x3F))); }
                    break;
                }
                default: out.push_back(e);
            }
        } else out.push_back(c);
    }
    if (p >= t.size()) throw JsonError("unterminated string");
    ++p;
    return out;
}

Json Json::parse(const std::string &t, size_t &p) {
    skip_ws(t, p);
    if (p >= t.size()) throw JsonError("unexpected end of JSON");
    char c = t[p];
    Json j;
    if (c == '{') {
        j.kind = Object;
        ++p;
        skip_ws(t, p);
        if (p < t.size() && t[p] == '}') { ++p; return j; }
        for (;;) {
            skip_ws(t, p);
            if (p >= t.size()) throw JsonError("unterminated object");
            std::string k = parse_string(t, p);
            skip_ws(t, p);
            if (p >= t.size() || t[p] != ':') throw JsonError("expected ':' at offset " + std::to_string(p));
            ++p;
            Json v = parse(t, p);
            j.obj.emplace_back(std::move(k), std::move(v));
            skip_ws(t, p);
            if (p < t.size() && t[p] == ',') { ++p; continue; }
            if (p < t.size() && t[p] == '}') { ++p; break; }
            throw JsonError("expected ',' or '}' at offset " + std::to_string(p));
        }
        return j;
    }
    if (c == '[') {
        j.kind = Array;
        ++p;
        skip_ws(t, p);
        if (p < t.size() && t[p] == ']') { ++p; return j; }
        for (;;) {
            j.arr.push_back(parse(t, p));
            skip_ws(t, p);
            if (p < t.size() && t[p] == ',') { ++p; continue; }
            if (p < t.size() && t[p] == ']') { ++p; break; }
            throw JsonError("expected ',' or ']' at offset " + std::to_string(p));
        }
        return j;
    }
    if (c == '"') { j.kind = String; j.s = parse_string(t, p); return j; }
    if (t.compare(p, 4, "true") == 0) { p += 4; return boolean(true); }
    if (t.compare(p, 5, "false") == 0) { p += 5; return boolean(false); }
    if (t.compare(p, 4, "null") == 0) { p += 4; return j; }
    size_t q = p;
    bool is_double = false;
    if (q < t.size() && (t[q] == '-' || t[q] == '+')) ++q;
    while (q < t.size() && (std::isdigit((unsigned char)t[q]) || t[q] == '.' || t[q] == 'e' || t[q] == 'E' || t[q] == '-' || t[q] == '+')) {
        if (t[q] == '.' || t[q] == 'e' || t[q] == 'E') is_double = true;
        ++q;
    }
    if (q == p) throw JsonError("unexpected character '" + std::string(1, c) + "' at offset " + std::to_string(p));
    std::string num = t.substr(p, q - p);
    p = q;
    if (is_double) { j.kind = Double; j.d = strtod(num.c_str(), nullptr); }
    else { j.kind = Int; j.i = strtoll(num.c_str(), nullptr, 10); }
    return j;
}

Json Json::parse(const std::string &text) {
    size_t p = 0;
    Json j = parse(text, p);
    skip_ws(text, p);
    if (p != text.size()) throw JsonError("trailing characters after JSON value at offset " + std::to_string(p));
    return j;
}

}  // namespace host
}  // namespace dissem
