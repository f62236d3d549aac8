// hj_json.h — a small JSON reader for the two JSON inputs of the host loop: render_option.json
// (/root/reference/include/loader/render_json_loader.h:97-202, read there with nlohmann::json) and the glTF 2.0
// scene (/root/reference/include/loader/gltfloader.h:1068, read there with tinygltf). Numbers are kept as double
// (what both libraries do) and narrowed by the caller exactly where the reference narrows them.
#pragma once
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace hj {

struct Json {
    enum Type { Null, Bool, Number, String, Array, Object } type = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;   // insertion order kept (glTF attribute order matters to nobody, but cheap)

    bool is_null() const { return type == Null; }
    bool has(const std::string& k) const { return find(k) != nullptr; }
    const Json* find(const std::string& k) const {
        if (type != Object) return nullptr;
        for (auto& kv : obj) if (kv.first == k) return &kv.second;
        return nullptr;
    }
    // nlohmann-like accessors: a missing key or a wrong type is an error (the reference's loader throws and returns false)
    const Json& at(const std::string& k) const {
        const Json* j = find(k);
        if (!j) throw std::runtime_error("json: missing key \"" + k + "\"");
        return *j;
    }
    const Json& at(size_t i) const {
        if (type != Array || i >= arr.size()) throw std::runtime_error("json: array index out of range");
        return arr[i];
    }
    size_t size() const { return type == Array ? arr.size() : (type == Object ? obj.size() : 0); }
    double number() const { if (type != Number) throw std::runtime_error("json: number expected"); return num; }
    bool boolean() const { if (type != Bool) throw std::runtime_error("json: bool expected"); return b; }
    const std::string& string() const { if (type != String) throw std::runtime_error("json: string expected"); return str; }
    // optional lookups with defaults (tinygltf fills defaults for absent glTF properties)
    double number_or(const std::string& k, double d) const { const Json* j = find(k); return (j && j->type == Number) ? j->num : d; }
    int int_or(const std::string& k, int d) const { const Json* j = find(k); return (j && j->type == Number) ? (int)j->num : d; }
    std::string string_or(const std::string& k, const std::string& d) const { const Json* j = find(k); return (j && j->type == String) ? j->str : d; }
};

class JsonParser {
public:
    static Json parse(const std::string& text) {
        JsonParser p(text);
        p.ws();
        Json j = p.value();
        p.ws();
        if (p.i != p.s.size()) p.fail("trailing characters");
        return j;
    }

private:
    const std::string& s;
    size_t i = 0;
    explicit JsonParser(const std::string& t) : s(t) {}
    [[noreturn]] void fail(const char* what) const { throw std::runtime_error(std::string("json: ") + what + " at offset " + std::to_string(i)); }
    void ws() { while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\n' || s[i] == '\r')) i++; }
    bool lit(const char* w) { size_t n = strlen(w); if (s.compare(i, n, w) == 0) { i += n; return true; } return false; }

    Json value() {
        if (i >= s.size()) fail("unexpected end");
        Json j;
        const char c = s[i];
        if (c == '{') {
            j.type = Json::Object; i++; ws();
            if (i < s.size() && s[i] == '}') { i++; return j; }
            while (true) {
                ws();
                if (i >= s.size() || s[i] != '"') fail("object key expected");
                std::string k = str();
                ws();
                if (i >= s.size() || s[i] != ':') fail("':' expected");
                i++; ws();
                j.obj.emplace_back(std::move(k), value());
                ws();
                if (i < s.size() && s[i] == ',') { i++; continue; }
                if (i < s.size() && s[i] == '}') { i++; break; }
                fail("',' or '}' expected");
            }
        } else if (c == '[') {
            j.type = Json::Array; i++; ws();
            if (i < s.size() && s[i] == ']') { i++; return j; }
            while (true) {
                ws();
                j.arr.push_back(value());
                ws();
                if (i < s.size() && s[i] == ',') { i++; continue; }
                if (i < s.size() && s[i] == ']') { i++; break; }
                fail("',' or ']' expected");
            }
        } else if (c == '"') {
            j.type = Json::String; j.str = str();
        } else if (lit("true")) { j.type = Json::Bool; j.b = true; }
        else if (lit("false")) { j.type = Json::Bool; j.b = false; }
        else if (lit("null")) { j.type = Json::Null; }
        else {
            const char* b = s.c_str() + i; char* e = nullptr;
            j.num = strtod(b, &e);
            if (e == b) fail("value expected");
            j.type = Json::Number; i += (size_t)(e - b);
        }
        return j;
    }

    static void utf8(std::string& o, unsigned cp) {
        if (cp < 0x80) o += (char)cp;
        else if (cp < 0x800) { o += (char)(0xC0 | (cp >> 6)); o += (char)(0x80 | (cp & 0x3F)); }
        else if (cp < 0x10000) { o += (char)(0xE0 | (cp >> 12)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
        else { o += (char)(0xF0 | (cp >> 18)); o += (char)(0x80 | ((cp >> 12) & 0x3F)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
    }
    unsigned hex4() {
        if (i + 4 > s.size()) fail("bad \\u escape");
        unsigned v = 0;
        for (int k = 0; k < 4; k++) {
            const char c = s[i++]; v <<= 4;
            if (c >= '0' && c <= '9') v |= (unsigned)(c - '0');
            else if (c >= 'a' && c <= 'f') v |= (unsigned)(c - 'a' + 10);
            else if (c >= 'A' && c <= 'F') v |= (unsigned)(c - 'A' + 10);
            else fail("bad \\u escape");
        }
        return v;
    }
    std::string str() {
        std::string o;
        i++;   // opening quote
        while (true) {
            if (i >= s.size()) fail("unterminated string");
            const char c = s[i++];
            if (c == '"') break;
            if (c != '\\') { o += c; continue; }
            if (i >= s.size()) fail("unterminated escape");
            const char e = s[i++];
            switch (e) {
                case '"': o += '"'; break; case '\\': o += '\\'; break; case '/': o += '/'; break;
                case 'b': o += '\b'; break; case 'f': o += '\f'; break; case 'n': o += '\n'; break;
                case 'r': o += '\r'; break; case 't': o += '\t'; break;
                case 'u': {
                    unsigned cp = hex4();
                    if (cp >= 0xD800 && cp < 0xDC00 && i + 1 < s.size() && s[i] == '\\' && s[i + 1] == 'u') {
                        i += 2; const unsigned lo = hex4(); cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                    }
                    utf8(o, cp); break;
                }
                default: fail("bad escape");
            }
        }
        return o;
    }
};

}  // namespace hj
