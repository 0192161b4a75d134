// Minimal JSON value (parse + serialise) for the config file and the control-plane envelope.
// Stands in for Go's encoding/json as the reference uses it (cmd/config.go:58-59, message.go:273-301).
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace dissem {
namespace host {

struct JsonError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

class Json {
public:
    enum Kind { Null, Bool, Int, Double, String, Array, Object };
    Kind kind = Null;
    bool b = false;
    int64_t i = 0;
    double d = 0.0;
    std::string s;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;   // insertion order kept (Go marshals struct fields in order)

    Json() = default;
    static Json of(int64_t v) { Json j; j.kind = Int; j.i = v; return j; }
    static Json of(const std::string &v) { Json j; j.kind = String; j.s = v; return j; }
    static Json boolean(bool v) { Json j; j.kind = Bool; j.b = v; return j; }
    static Json object() { Json j; j.kind = Object; return j; }
    static Json array() { Json j; j.kind = Array; return j; }

    Json &set(const std::string &k, Json v) { obj.emplace_back(k, std::move(v)); return *this; }
    Json &push(Json v) { arr.push_back(std::move(v)); return *this; }

    bool is_object() const { return kind == Object; }
    bool is_number() const { return kind == Int || kind == Double; }
    int64_t as_int() const {
        if (kind == Int) return i;
        if (kind == Double) return (int64_t)d;
        if (kind == Bool) return b;
        return 0;
    }
    // exact key, then case-insensitive (Go's decoder: `Id` fills field `ID`)
    const Json *get(const std::string &key) const;
    std::string dump() const;

    // Parses ONE value starting at text[pos]; pos is advanced past it (streams of concatenated objects, as
    // on the reference's TCP connections, are read by calling this repeatedly).
    static Json parse(const std::string &text, size_t &pos);
    static Json parse(const std::string &text);
};

}  // namespace host
}  // namespace dissem
