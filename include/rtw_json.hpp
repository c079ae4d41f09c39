// rtw_json.hpp — the reference's JSON scene wire format for C++ callers of rtw.hpp: the
// `Into<JsonValue>` / `TryFrom<JsonValue>` impls of the Rust tree, restated (the Python host side has
// the same in raytracing-in-a-weekend_b200/scene_json.py, and the two read each other's text).
//
//     Vec3          {"x":..,"y":..,"z":..}                                   vec3.rs:22-50
//     Material      {"metallicness":..,"opacity":..,"ir":..}                 materials.rs:21-54
//                   (`emmited` is NOT serialised; parsing sets it to 0)
//     ImageTexture  {"row":W,"col":H,"img":[Vec3,...]}                       texture.rs:28-59,268-276
//                   (noise is NOT serialised; parsing sets noise = None, noise_scale = 1)
//     Sphere        {"origin","radius","col_mod","material","velocity","texture"}   sphere.rs:44-90
//     Scene         {"spheres":[Sphere,...]}  -> Scene::new_sphere            viewport.rs:174-205
//                   (quads and instances are NOT serialised)
//
// Errors follow the reference: any missing or mistyped field is a ParseError (viewport.rs:2-26), never a
// partial scene. Numbers are f32 on both sides (`as_f32`); they are written with nine significant
// digits, which round-trips every f32. The parser is the subset of JSON the `json` crate (0.12.4,
// Rust/Cargo.lock) emits and more: objects, arrays, numbers, strings with escapes, true/false/null.
#ifndef RTW_JSON_HPP
#define RTW_JSON_HPP

#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

#include "rtw.hpp"

namespace rtw {
namespace json {

struct ParseError : std::runtime_error {   // viewport.rs:2-26 errors::ParseError { source }
    explicit ParseError(const std::string& source = "None") : std::runtime_error("ParseError: " + source) {}
};

struct Value {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Value> arr;
    std::vector<std::pair<std::string, Value>> obj;   // insertion order, as the json crate keeps it

    const Value* get(const char* key) const {         // JsonValue's `value[key]`: Null when absent
        if (kind == Object) for (const auto& kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
    static Value number(double x) { Value v; v.kind = Number; v.num = x; return v; }
    static Value object() { Value v; v.kind = Object; return v; }
    static Value array() { Value v; v.kind = Array; return v; }
    Value& set(const char* key, Value v) { obj.emplace_back(key, std::move(v)); return *this; }
};

// ----------------------------------------------------------------------------------------- text
class Parser {
public:
    explicit Parser(const std::string& s) : s_(s) {}
    Value parse() {
        Value v = value();
        ws();
        if (i_ != s_.size()) fail("trailing characters");
        return v;
    }
private:
    [[noreturn]] void fail(const char* what) const { throw ParseError(std::string(what) + " at byte " + std::to_string(i_)); }
    void ws() { while (i_ < s_.size() && (s_[i_] == ' ' || s_[i_] == '\n' || s_[i_] == '\t' || s_[i_] == '\r')) ++i_; }
    bool eat(char c) { ws(); if (i_ < s_.size() && s_[i_] == c) { ++i_; return true; } return false; }
    void word(const char* w) { for (; *w; ++w) { if (i_ >= s_.size() || s_[i_] != *w) fail("bad literal"); ++i_; } }
    Value value() {
        if (++depth_ > 512) fail("nesting too deep");
        ws();
        if (i_ >= s_.size()) fail("unexpected end");
        Value v;
        const char c = s_[i_];
        if (c == '{') {
            ++i_; v.kind = Value::Object;
            if (!eat('}')) {
                do {
                    ws();
                    if (i_ >= s_.size() || s_[i_] != '"') fail("object key expected");
                    std::string k = string();
                    if (!eat(':')) fail("':' expected");
                    v.obj.emplace_back(std::move(k), value());
                } while (eat(','));
                if (!eat('}')) fail("'}' expected");
            }
        } else if (c == '[') {
            ++i_; v.kind = Value::Array;
            if (!eat(']')) {
                do v.arr.push_back(value()); while (eat(','));
                if (!eat(']')) fail("']' expected");
            }
        } else if (c == '"') { v.kind = Value::String; v.str = string(); }
        else if (c == 't') { word("true"); v.kind = Value::Bool; v.b = true; }
        else if (c == 'f') { word("false"); v.kind = Value::Bool; v.b = false; }
        else if (c == 'n') { word("null"); }
        else {
            const char* begin = s_.c_str() + i_;
            char* end = nullptr;
            if (!(c == '-' || (c >= '0' && c <= '9'))) fail("value expected");
            v.kind = Value::Number; v.num = std::strtod(begin, &end);
            if (end == begin) fail("bad number");
            i_ += (size_t)(end - begin);
        }
        --depth_;
        return v;
    }
    std::string string() {
        std::string out;
        ++i_;   // opening quote
        for (;;) {
            if (i_ >= s_.size()) fail("unterminated string");
            char c = s_[i_++];
            if (c == '"') return out;
            if (c != '\\') { out.push_back(c); continue; }
            if (i_ >= s_.size()) fail("unterminated escape");
            c = s_[i_++];
            switch (c) {
                case 'n': out.push_back('\n'); break; case 't': out.push_back('\t'); break;
                case 'r': out.push_back('\r'); break; case 'b': out.push_back('\b'); break;
                case 'f': out.push_back('\f'); break;
                case 'u': {
                    if (i_ + 4 > s_.size()) fail("bad \\u escape");
                    unsigned cp = (unsigned)std::strtoul(s_.substr(i_, 4).c_str(), nullptr, 16); i_ += 4;
                    if (cp < 0x80) out.push_back((char)cp);
                    else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 0x3F))); }
                    else { out.push_back((char)(0xE0 | (cp >> 12))); out.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out.push_back((char)(0x80 | (cp & 0x3F))); }
                    break;
                }
                default: out.push_back(c);   // \" \\ \/
            }
        }
    }
    const std::string& s_;
    size_t i_ = 0;
    int depth_ = 0;
};

inline void write(const Value& v, std::string& out) {
    switch (v.kind) {
        case Value::Null: out += "null"; break;
        case Value::Bool: out += v.b ? "true" : "false"; break;
        case Value::Number: {
            char buf[40];
            if (v.num == (double)(long long)v.num && std::fabs(v.num) < 1e15) std::snprintf(buf, sizeof buf, "%lld", (long long)v.num);
            else std::snprintf(buf, sizeof buf, "%.9g", v.num);
            out += buf; break;
        }
        case Value::String: {
            out.push_back('"');
            for (char c : v.str) {
                if (c == '"' || c == '\\') { out.push_back('\\'); out.push_back(c); }
                else if (c == '\n') out += "\\n";
                else if ((unsigned char)c < 0x20) { char b[8]; std::snprintf(b, sizeof b, "\\u%04x", c); out += b; }
                else out.push_back(c);
            }
            out.push_back('"'); break;
        }
        case Value::Array:
            out.push_back('[');
            for (size_t i = 0; i < v.arr.size(); ++i) { if (i) out.push_back(','); write(v.arr[i], out); }
            out.push_back(']'); break;
        case Value::Object:
            out.push_back('{');
            for (size_t i = 0; i < v.obj.size(); ++i) {
                if (i) out.push_back(',');
                Value k; k.kind = Value::String; k.str = v.obj[i].first;
                write(k, out); out.push_back(':'); write(v.obj[i].second, out);
            }
            out.push_back('}'); break;
    }
}

// ------------------------------------------------------------------------------- to JSON values
inline Value to_json(Vec3 v) {                                   // vec3.rs:22-30
    return Value::object().set("x", Value::number(v.x)).set("y", Value::number(v.y)).set("z", Value::number(v.z));
}
inline Value to_json(const Material& m) {                        // materials.rs:21-29
    return Value::object().set("metallicness", Value::number(m.metallicness)).set("opacity", Value::number(m.opacity)).set("ir", Value::number(m.ir));
}
inline Value to_json(const ImageTexture& t) {                    // texture.rs:268-276
    Value img = Value::array();
    for (size_t i = 0; i + 2 < t.img.size(); i += 3) img.arr.push_back(to_json(Vec3(t.img[i], t.img[i + 1], t.img[i + 2])));
    return Value::object().set("row", Value::number(t.row)).set("col", Value::number(t.col)).set("img", std::move(img));
}
inline Value to_json(const Sphere& s) {                          // sphere.rs:44-55
    return Value::object().set("origin", to_json(s.origin)).set("radius", Value::number(s.radius)).set("col_mod", to_json(s.col_mod))
        .set("material", to_json(s.mat)).set("velocity", to_json(s.velocity)).set("texture", to_json(*s.texture));
}
inline Value to_json(const Scene& sc) {                          // viewport.rs:174-180
    Value sp = Value::array();
    for (const Sphere& s : sc.spheres) sp.arr.push_back(to_json(s));
    return Value::object().set("spheres", std::move(sp));
}

// ----------------------------------------------------------------------------- from JSON values
inline float as_f32(const Value* parent, const char* key) {      // JsonValue::as_f32 on value[key]: a number or nothing
    const Value* v = parent ? parent->get(key) : nullptr;
    if (!v || v->kind != Value::Number) throw ParseError();
    return (float)v->num;
}
inline uint32_t as_usize(const Value* parent, const char* key) {
    const Value* v = parent ? parent->get(key) : nullptr;
    if (!v || v->kind != Value::Number || v->num < 0 || v->num != (double)(uint64_t)v->num || v->num > 4294967295.0) throw ParseError();
    return (uint32_t)v->num;
}
inline Vec3 vec3_from_json(const Value* v) { return Vec3(as_f32(v, "x"), as_f32(v, "y"), as_f32(v, "z")); }   // vec3.rs:31-50
inline Material material_from_json(const Value* v) {             // materials.rs:30-54 (emmited = 0)
    return Material{as_f32(v, "metallicness"), as_f32(v, "opacity"), as_f32(v, "ir"), Vec3()};
}
inline ImageTexture::Ptr texture_from_json(const Value* v) {     // texture.rs:28-59
    const Value* img = v ? v->get("img") : nullptr;
    if (!img || img->kind != Value::Array) throw ParseError();
    std::vector<float> texels;
    texels.reserve(img->arr.size() * 3);
    for (const Value& p : img->arr) { const Vec3 c = vec3_from_json(&p); texels.push_back(c.x); texels.push_back(c.y); texels.push_back(c.z); }
    const uint32_t row = as_usize(v, "row"), col = as_usize(v, "col");
    // the reference accepts row*col != len(img) and panics at the first texel fetch past the end; here it is a ParseError
    if ((uint64_t)row * col * 3 != texels.size()) throw ParseError("img must hold row*col texels");
    return ImageTexture::make(std::move(texels), row, col, nullptr, 1.0f);
}
inline Sphere sphere_from_json(const Value* v) {                 // sphere.rs:56-90
    if (!v || v->kind != Value::Object) throw ParseError();
    return Sphere{vec3_from_json(v->get("origin")), as_f32(v, "radius"), vec3_from_json(v->get("col_mod")),
                  material_from_json(v->get("material")), vec3_from_json(v->get("velocity")), texture_from_json(v->get("texture"))};
}
inline Scene scene_from_json(const Value& v) {                   // viewport.rs:181-205
    const Value* sp = v.get("spheres");
    if (!sp || sp->kind != Value::Array) throw ParseError();
    std::vector<Sphere> spheres;
    for (const Value& s : sp->arr) spheres.push_back(sphere_from_json(&s));
    return Scene::new_sphere(std::move(spheres));
}

// `json::stringify(scene.into())`
inline std::string dumps(const Scene& scene) { std::string out; write(to_json(scene), out); return out; }
// `Scene::try_from(json::parse(text)?)`
inline Scene loads(const std::string& text) { return scene_from_json(Parser(text).parse()); }

}  // namespace json
}  // namespace rtw
#endif  // RTW_JSON_HPP
