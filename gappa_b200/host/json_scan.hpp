// Minimal pull scanner over an in-memory JSON text — just what a jplace document needs.
//
// No DOM is built (the reference parses each file into a full JsonDocument first,
// genesis utils/formats/json/reader.cpp:55-343; at 100k placements per sample that is the ingest
// bottleneck).  Numbers are accumulated digit by digit exactly the way the
// reference does (reader.cpp:246-343: integer part in a uint64, fraction as dx += digit * 10^-k
// with the power formed by repeated division, exponent by square-and-multiply), because that
// routine is not correctly rounded and a strtod-based parser would differ from it by an ulp.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>

namespace krdhost {

struct JsonNumber {
    enum Kind : uint8_t { Unsigned, Signed, Float } kind = Unsigned;
    uint64_t u = 0;   // magnitude for Unsigned; two's complement value for Signed
    double d = 0.0;   // Float

    double as_double() const
    {
        if (kind == Float) return d;
        if (kind == Unsigned) return (double)u;
        return (double)(int64_t)u;
    }
    uint64_t as_size() const
    {
        if (kind == Float) return (uint64_t)d;
        return u;   // Signed: wraps like the reference's static_cast
    }
};

class JsonScanner {
public:
    JsonScanner(const char* begin, const char* end, std::string source) : p_(begin), b_(begin), e_(end), src_(std::move(source)) {}

    void ws() { while (p_ < e_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) ++p_; }
    bool at_end() { ws(); return p_ >= e_; }
    char peek() { ws(); return p_ < e_ ? *p_ : '\0'; }
    bool try_char(char c) { ws(); if (p_ < e_ && *p_ == c) { ++p_; return true; } return false; }
    void expect(char c)
    {
        if (!try_char(c)) fail(std::string("expecting '") + c + "'");
    }

    [[noreturn]] void fail(const std::string& what) const
    {
        throw std::runtime_error("Malformed JSON in " + src_ + " at byte " + std::to_string(p_ - b_) + ": " + what + ".");
    }

    enum Type { Object, Array, String, Number, True, False, Null, Bad };
    Type type()
    {
        switch (peek()) {
            case '{': return Object;
            case '[': return Array;
            case '"': return String;
            case 't': return True;
            case 'f': return False;
            case 'n': return Null;
            case '-': case '+': case '.': return Number;
            default: return (p_ < e_ && *p_ >= '0' && *p_ <= '9') ? Number : Bad;
        }
    }
    static const char* type_name(Type t)
    {
        switch (t) {
            case Object: return "object"; case Array: return "array"; case String: return "string";
            case Number: return "number"; case True: case False: return "boolean"; case Null: return "null";
            default: return "invalid";
        }
    }

    std::string string()
    {
        expect('"');
        std::string out;
        for (;;) {
            if (p_ >= e_) fail("unterminated string");
            char c = *p_++;
            if (c == '"') return out;
            if (c != '\\') { out += c; continue; }
            if (p_ >= e_) fail("unterminated escape");
            c = *p_++;
            switch (c) {
                case '"': out += '"'; break;
                case '\\': out += '\\'; break;
                case '/': out += '/'; break;
                case 'b': out += '\b'; break;
                case 'f': out += '\f'; break;
                case 'n': out += '\n'; break;
                case 'r': out += '\r'; break;
                case 't': out += '\t'; break;
                case 'u': {
                    if (e_ - p_ < 4) fail("short \\u escape");
                    unsigned cp = 0;
                    for (int k = 0; k < 4; ++k) {
                        char h = *p_++;
                        cp <<= 4;
                        if (h >= '0' && h <= '9') cp |= (unsigned)(h - '0');
                        else if (h >= 'a' && h <= 'f') cp |= (unsigned)(h - 'a' + 10);
                        else if (h >= 'A' && h <= 'F') cp |= (unsigned)(h - 'A' + 10);
                        else fail("bad \\u escape");
                    }
                    if (cp < 0x80) out += (char)cp;
                    else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
                    else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
                    break;
                }
                default: fail("invalid escape");
            }
        }
    }

    void skip_string()
    {
        expect('"');
        while (p_ < e_) {
            char c = *p_++;
            if (c == '"') return;
            if (c == '\\') { if (p_ < e_) ++p_; }
        }
        fail("unterminated string");
    }

    JsonNumber number()
    {
        ws();
        if (p_ >= e_) fail("expecting number");
        JsonNumber n;
        bool neg = false;
        if (*p_ == '-') { neg = true; ++p_; } else if (*p_ == '+') ++p_;
        bool mant = false;
        uint64_t ix = 0;
        while (p_ < e_ && *p_ >= '0' && *p_ <= '9') { ix = ix * 10 + (uint64_t)(*p_ - '0'); ++p_; mant = true; }
        const char c = p_ < e_ ? *p_ : '\0';
        if (mant && !(c == '.' || c == 'e' || c == 'E')) {
            if (neg) { n.kind = JsonNumber::Signed; n.u = (uint64_t)0 - ix; }
            else { n.kind = JsonNumber::Unsigned; n.u = ix; }
            return n;
        }
        double dx = (double)ix;
        if (p_ < e_ && *p_ == '.') {
            ++p_;
            // the reference forms the place value by repeated division (pos /= 10.0 per digit, reader.cpp:290-296); that
            // sequence does not depend on the input, so it is tabulated ONCE with the same divisions — same bits, without a
            // serial chain of ~17 divides per number (64 -> 40 ns per number on 17-digit fractions)
            const double* const pw = frac_place_values();
            int k = 0;
            double pos = 1.0;
            while (p_ < e_ && *p_ >= '0' && *p_ <= '9') {
                if (k < FRAC_TABLE) pos = pw[k++]; else pos /= 10.0;
                dx += (double)(*p_ - '0') * pos;
                ++p_; mant = true;
            }
        }
        if (!mant) fail("invalid float number");
        if (p_ < e_ && (*p_ == 'e' || *p_ == 'E')) {
            ++p_;
            bool eneg = false;
            if (p_ < e_ && (*p_ == '-' || *p_ == '+')) { eneg = (*p_ == '-'); ++p_; }
            if (!(p_ < e_ && *p_ >= '0' && *p_ <= '9')) fail("expecting digits in exponent");
            long ex = 0;
            while (p_ < e_ && *p_ >= '0' && *p_ <= '9') { if (ex < 100000) ex = ex * 10 + (*p_ - '0'); ++p_; }
            if (ex != 0) {
                double base = eneg ? 0.1 : 10.0;
                while (ex != 1) {
                    if ((ex & 1) == 0) { base = base * base; ex >>= 1; }
                    else { dx *= base; --ex; }
                }
                dx *= base;
            }
        }
        n.kind = JsonNumber::Float;
        n.d = neg ? -dx : dx;
        return n;
    }

    void skip_value()
    {
        switch (type()) {
            case Object:
                ++p_;
                if (try_char('}')) return;
                for (;;) {
                    skip_string(); expect(':'); skip_value();
                    if (try_char(',')) continue;
                    expect('}');
                    return;
                }
            case Array:
                ++p_;
                if (try_char(']')) return;
                for (;;) {
                    skip_value();
                    if (try_char(',')) continue;
                    expect(']');
                    return;
                }
            case String: skip_string(); return;
            case Number: number(); return;
            case True: literal("true"); return;
            case False: literal("false"); return;
            case Null: literal("null"); return;
            default: fail("unexpected character");
        }
    }

    void literal(const char* w)
    {
        ws();
        for (const char* q = w; *q; ++q) { if (p_ >= e_ || *p_ != *q) fail(std::string("expecting '") + w + "'"); ++p_; }
    }

    const std::string& source() const { return src_; }

    static constexpr int FRAC_TABLE = 40;
    static const double* frac_place_values()       // [k] = 1.0 divided k + 1 times by 10.0, one division after the other
    {
        static const struct Table {
            double v[FRAC_TABLE];
            Table() { double pos = 1.0; for (int k = 0; k < FRAC_TABLE; ++k) { pos /= 10.0; v[k] = pos; } }
        } t;
        return t.v;
    }

private:
    const char* p_;
    const char* b_;
    const char* e_;
    std::string src_;
};

}  // namespace krdhost
