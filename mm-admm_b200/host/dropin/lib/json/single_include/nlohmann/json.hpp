// Minimal JSON value under the nlohmann::json name, sufficient for the MM-ADMM driver (main.cpp:2,801-867):
// the experiment files in Experiments/InputFiles/*.json are flat objects of numbers, booleans and strings.
// nlohmann/json is git-ignored in the reference tree (lib/json), so the drop-in build supplies
// this header at the path main.cpp includes.  Supported: `stream >> j`, j["key"] (read, assign), implicit
// conversion to bool / integer / floating / std::string, contains(), `stream << j` (compact, keys sorted --
// the format nlohmann's dump() prints), nested objects and arrays (parsed and printed, indexable by position).
#ifndef MMADMM_JSON_SHIM_HPP
#define MMADMM_JSON_SHIM_HPP
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <iterator>
#include <map>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

namespace nlohmann {

class json {
public:
    enum class Kind { Null, Bool, Int, Float, String, Object, Array };

    json() : kind_(Kind::Null) {}
    json(bool b) : kind_(Kind::Bool), i_(b) {}
    json(int v) : kind_(Kind::Int), i_(v) {}
    json(long v) : kind_(Kind::Int), i_(v) {}
    json(long long v) : kind_(Kind::Int), i_(v) {}
    json(unsigned v) : kind_(Kind::Int), i_(v) {}
    json(double v) : kind_(Kind::Float), d_(v) {}
    json(const char *s) : kind_(Kind::String), s_(s) {}
    json(const std::string &s) : kind_(Kind::String), s_(s) {}

    Kind kind() const { return kind_; }
    bool is_null() const { return kind_ == Kind::Null; }
    bool is_object() const { return kind_ == Kind::Object; }
    bool is_number() const { return kind_ == Kind::Int || kind_ == Kind::Float; }
    size_t size() const { return kind_ == Kind::Object ? obj_.size() : kind_ == Kind::Array ? arr_.size() : (kind_ == Kind::Null ? 0 : 1); }

    // operator[] on a null value turns it into an object, like nlohmann's
    json &operator[](const std::string &key) {
        if (kind_ == Kind::Null) kind_ = Kind::Object;
        if (kind_ != Kind::Object) throw std::domain_error("json: operator[] with a string key on a non-object");
        return obj_[key];
    }
    json &operator[](const char *key) { return (*this)[std::string(key)]; }
    const json &at(const std::string &key) const {
        auto it = obj_.find(key);
        if (kind_ != Kind::Object || it == obj_.end()) throw std::out_of_range("json: key '" + key + "' not found");
        return it->second;
    }
    json &operator[](size_t i) { if (kind_ != Kind::Array) throw std::domain_error("json: not an array"); return arr_.at(i); }
    bool contains(const std::string &key) const { return kind_ == Kind::Object && obj_.count(key) != 0; }

    // implicit conversions (never to char, so that `std::string s; s = j;` stays unambiguous)
    template <typename T, typename = std::enable_if_t<(std::is_arithmetic<T>::value && !std::is_same<T, char>::value) ||
                                                      std::is_same<T, std::string>::value>>
    operator T() const { return get<T>(); }

    template <typename T> T get() const {
        if constexpr (std::is_same<T, std::string>::value) {
            if (kind_ != Kind::String) throw std::domain_error("json: value is not a string");
            return s_;
        } else if constexpr (std::is_same<T, bool>::value) {
            if (kind_ == Kind::Bool || kind_ == Kind::Int) return i_ != 0;
            throw std::domain_error("json: value is not a boolean");
        } else {
            if (kind_ == Kind::Int || kind_ == Kind::Bool) return static_cast<T>(i_);
            if (kind_ == Kind::Float) return static_cast<T>(d_);
            throw std::domain_error("json: value is not a number");
        }
    }

    std::string dump() const { std::string out; write(out); return out; }

    static json parse(const std::string &text) {
        size_t pos = 0;
        json j = parseValue(text, pos);
        skipWs(text, pos);
        if (pos != text.size()) throw std::invalid_argument("json: trailing characters at offset " + std::to_string(pos));
        return j;
    }

private:
    Kind kind_;
    long long i_ = 0;
    double d_ = 0;
    std::string s_;
    std::map<std::string, json> obj_;
    std::vector<json> arr_;

    static void skipWs(const std::string &t, size_t &p) { while (p < t.size() && std::isspace((unsigned char)t[p])) p++; }
    static std::string parseString(const std::string &t, size_t &p) {
        std::string out;
        p++;  // opening quote
        while (p < t.size() && t[p] != '"') {
            char c = t[p++];
            if (c == '\\' && p < t.size()) {
                char e = t[p++];
                switch (e) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    default: out += e;  // \" \\ \/ (and \uXXXX is not needed by the input files)
                }
            } else {
                out += c;
            }
        }
        if (p >= t.size()) throw std::invalid_argument("json: unterminated string");
        p++;
        return out;
    }
    static json parseValue(const std::string &t, size_t &p) {
        skipWs(t, p);
        if (p >= t.size()) throw std::invalid_argument("json: unexpected end of input");
        char c = t[p];
        if (c == '{') {
            json j; j.kind_ = Kind::Object; p++;
            skipWs(t, p);
            if (p < t.size() && t[p] == '}') { p++; return j; }
            for (;;) {
                skipWs(t, p);
                if (p >= t.size() || t[p] != '"') throw std::invalid_argument("json: expected a key at offset " + std::to_string(p));
                std::string key = parseString(t, p);
                skipWs(t, p);
                if (p >= t.size() || t[p] != ':') throw std::invalid_argument("json: expected ':' at offset " + std::to_string(p));
                p++;
                j.obj_[key] = parseValue(t, p);
                skipWs(t, p);
                if (p < t.size() && t[p] == ',') { p++; continue; }
                if (p < t.size() && t[p] == '}') { p++; return j; }
                throw std::invalid_argument("json: expected ',' or '}' at offset " + std::to_string(p));
            }
        }
        if (c == '[') {
            json j; j.kind_ = Kind::Array; p++;
            skipWs(t, p);
            if (p < t.size() && t[p] == ']') { p++; return j; }
            for (;;) {
                j.arr_.push_back(parseValue(t, p));
                skipWs(t, p);
                if (p < t.size() && t[p] == ',') { p++; continue; }
                if (p < t.size() && t[p] == ']') { p++; return j; }
                throw std::invalid_argument("json: expected ',' or ']' at offset " + std::to_string(p));
            }
        }
        if (c == '"') return json(parseString(t, p));
        if (t.compare(p, 4, "true") == 0) { p += 4; return json(true); }
        if (t.compare(p, 5, "false") == 0) { p += 5; return json(false); }
        if (t.compare(p, 4, "null") == 0) { p += 4; return json(); }
        // number: integer unless it has a fraction or an exponent
        size_t q = p;
        bool isFloat = false;
        if (q < t.size() && (t[q] == '-' || t[q] == '+')) q++;
        while (q < t.size() && (std::isdigit((unsigned char)t[q]) || t[q] == '.' || t[q] == 'e' || t[q] == 'E' || t[q] == '-' || t[q] == '+')) {
            if (t[q] == '.' || t[q] == 'e' || t[q] == 'E') isFloat = true;
            q++;
        }
        if (q == p) throw std::invalid_argument("json: unexpected character at offset " + std::to_string(p));
        std::string num = t.substr(p, q - p);
        p = q;
        if (isFloat) return json(std::strtod(num.c_str(), nullptr));
        return json(std::strtoll(num.c_str(), nullptr, 10));
    }
    static void writeString(std::string &out, const std::string &s) {
        out += '"';
        for (char c : s) {
            if (c == '"' || c == '\\') { out += '\\'; out += c; }
            else if (c == '\n') out += "\\n";
            else if (c == '\t') out += "\\t";
            else out += c;
        }
        out += '"';
    }
    void write(std::string &out) const {
        switch (kind_) {
            case Kind::Null: out += "null"; break;
            case Kind::Bool: out += i_ ? "true" : "false"; break;
            case Kind::Int: out += std::to_string(i_); break;
            case Kind::Float: {
                // shortest representation that round-trips
                char buf[40];
                for (int prec = 1; prec <= 17; prec++) {
                    std::snprintf(buf, sizeof buf, "%.*g", prec, d_);
                    if (std::strtod(buf, nullptr) == d_) break;
                }
                std::string s(buf);
                if (s.find_first_of(".eEn") == std::string::npos) s += ".0";
                out += s;
                break;
            }
            case Kind::String: writeString(out, s_); break;
            case Kind::Object: {
                out += '{';
                bool first = true;
                for (const auto &kv : obj_) {
                    if (!first) out += ',';
                    first = false;
                    writeString(out, kv.first);
                    out += ':';
                    kv.second.write(out);
                }
                out += '}';
                break;
            }
            case Kind::Array: {
                out += '[';
                for (size_t i = 0; i < arr_.size(); i++) { if (i) out += ','; arr_[i].write(out); }
                out += ']';
                break;
            }
        }
    }
};

inline std::istream &operator>>(std::istream &in, json &j) {
    std::string text((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
    j = json::parse(text);
    return in;
}
inline std::ostream &operator<<(std::ostream &out, const json &j) { return out << j.dump(); }

}  // namespace nlohmann
#endif
