// Json.hpp — minimal recursive-descent JSON reader for the glTF loader (objects, arrays, strings, numbers, bools, null).
#pragma once

#include <cstdlib>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace vkrt
{
struct JsonValue
{
    enum Type
    {
        Null,
        Bool,
        Number,
        String,
        Array,
        Object
    } type = Null;
    bool boolean = false;
    double number = 0.0;
    std::string string;
    std::vector<JsonValue> array;
    std::map<std::string, JsonValue> object;

    bool has(const std::string& key) const { return type == Object && object.count(key) != 0; }
    const JsonValue& operator[](const std::string& key) const
    {
        static const JsonValue null;
        const auto it = object.find(key);
        return it == object.end() ? null : it->second;
    }
    const JsonValue& operator[](size_t i) const
    {
        static const JsonValue null;
        return i < array.size() ? array[i] : null;
    }
    size_t size() const { return type == Array ? array.size() : (type == Object ? object.size() : 0); }
    long long asInt(long long fallback = 0) const { return type == Number ? (long long)number : fallback; }
    const std::string& asString() const { return string; }
};

class JsonParser
{
public:
    // Returns false on malformed input; `error` holds a short description.
    static bool parse(const std::string& text, JsonValue& out, std::string& error)
    {
        JsonParser p(text);
        p.skip();
        if (!p.value(out))
        {
            error = p.m_error.empty() ? "malformed JSON" : p.m_error;
            return false;
        }
        return true;
    }

private:
    explicit JsonParser(const std::string& t) : m_text(t) {}
    const std::string& m_text;
    size_t m_pos = 0;
    std::string m_error;

    void skip()
    {
        while (m_pos < m_text.size() && (m_text[m_pos] == ' ' || m_text[m_pos] == '\n' || m_text[m_pos] == '\r' || m_text[m_pos] == '\t'))
        {
            ++m_pos;
        }
    }
    bool fail(const char* what)
    {
        m_error = std::string(what) + " at offset " + std::to_string(m_pos);
        return false;
    }
    bool value(JsonValue& v)
    {
        skip();
        if (m_pos >= m_text.size())
        {
            return fail("unexpected end");
        }
        const char c = m_text[m_pos];
        if (c == '{')
        {
            v.type = JsonValue::Object;
            ++m_pos;
            skip();
            if (m_pos < m_text.size() && m_text[m_pos] == '}')
            {
                ++m_pos;
                return true;
            }
            for (;;)
            {
                skip();
                std::string key;
                if (!str(key))
                {
                    return false;
                }
                skip();
                if (m_pos >= m_text.size() || m_text[m_pos] != ':')
                {
                    return fail("expected ':'");
                }
                ++m_pos;
                if (!value(v.object[key]))
                {
                    return false;
                }
                skip();
                if (m_pos < m_text.size() && m_text[m_pos] == ',')
                {
                    ++m_pos;
                    continue;
                }
                if (m_pos < m_text.size() && m_text[m_pos] == '}')
                {
                    ++m_pos;
                    return true;
                }
                return fail("expected ',' or '}'");
            }
        }
        if (c == '[')
        {
            v.type = JsonValue::Array;
            ++m_pos;
            skip();
            if (m_pos < m_text.size() && m_text[m_pos] == ']')
            {
                ++m_pos;
                return true;
            }
            for (;;)
            {
                v.array.emplace_back();
                if (!value(v.array.back()))
                {
                    return false;
                }
                skip();
                if (m_pos < m_text.size() && m_text[m_pos] == ',')
                {
                    ++m_pos;
                    continue;
                }
                if (m_pos < m_text.size() && m_text[m_pos] == ']')
                {
                    ++m_pos;
                    return true;
                }
                return fail("expected ',' or ']'");
            }
        }
        if (c == '"')
        {
            v.type = JsonValue::String;
            return str(v.string);
        }
        if (m_text.compare(m_pos, 4, "true") == 0)
        {
            v.type = JsonValue::Bool, v.boolean = true, m_pos += 4;
            return true;
        }
        if (m_text.compare(m_pos, 5, "false") == 0)
        {
            v.type = JsonValue::Bool, v.boolean = false, m_pos += 5;
            return true;
        }
        if (m_text.compare(m_pos, 4, "null") == 0)
        {
            v.type = JsonValue::Null, m_pos += 4;
            return true;
        }
        char* end = nullptr;
        const double d = std::strtod(m_text.c_str() + m_pos, &end);
        if (end == m_text.c_str() + m_pos)
        {
            return fail("unexpected character");
        }
        v.type = JsonValue::Number;
        v.number = d;
        m_pos = (size_t)(end - m_text.c_str());
        return true;
    }
    bool str(std::string& out)
    {
        if (m_pos >= m_text.size() || m_text[m_pos] != '"')
        {
            return fail("expected string");
        }
        ++m_pos;
        while (m_pos < m_text.size() && m_text[m_pos] != '"')
        {
            char c = m_text[m_pos++];
            if (c == '\\' && m_pos < m_text.size())
            {
                const char e = m_text[m_pos++];
                switch (e)
                {
                case 'n': c = '\n'; break;
                case 't': c = '\t'; break;
                case 'r': c = '\r'; break;
                case 'b': c = '\b'; break;
                case 'f': c = '\f'; break;
                case 'u':
                {
                    const unsigned code = (unsigned)std::strtoul(m_text.substr(m_pos, 4).c_str(), nullptr, 16);
                    m_pos += 4;
                    if (code < 0x80)
                    {
                        c = (char)code;
                    }
                    else
                    {
                        out += (char)(0xC0 | (code >> 6 & 0x1F)); // BMP subset, 2-byte form is enough for file names
                        c = (char)(0x80 | (code & 0x3F));
                    }
                    break;
                }
                default: c = e; break;
                }
            }
            out += c;
        }
        if (m_pos >= m_text.size())
        {
            return fail("unterminated string");
        }
        ++m_pos;
        return true;
    }
};
} // namespace vkrt
