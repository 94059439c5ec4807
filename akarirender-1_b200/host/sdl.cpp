// sdl.cpp — see sdl.h.
#include "sdl.h"

#include <ctype.h>
#include <stdio.h>
#include <string.h>

#include <fstream>
#include <iterator>

namespace akb_host {

std::string dir_of(const std::string &path) {
    const size_t k = path.find_last_of('/');
    if (k == std::string::npos) return ".";
    return k == 0 ? "/" : path.substr(0, k);
}

std::string join_path(const std::string &dir, const std::string &rel) {
    if (!rel.empty() && rel[0] == '/') return rel;
    if (dir.empty() || dir == ".") return rel;
    return dir.back() == '/' ? dir + rel : dir + "/" + rel;
}

namespace {

struct Cursor {
    const std::string &src;
    const std::string &file;
    size_t pos = 0;
    int line = 1, col = 1;
    Module *mod;
    bool (*known)(const std::string &);

    char peek() const { return pos < src.size() ? src[pos] : '\0'; }
    char take() {
        const char c = peek();
        if (c == '\0') return c;
        pos++;
        if (c == '\n')
            line++, col = 1;
        else
            col++;
        return c;
    }
    bool at(const char *word) const { return src.compare(pos, strlen(word), word) == 0; }
    [[noreturn]] void fail(const std::string &msg) const {
        char where[64];
        snprintf(where, sizeof where, ":%d:%d: ", line, col);
        throw ParseError(file + where + msg);
    }
    void want(char c) {
        if (peek() != c) fail(std::string("'") + c + "' expected, found '" + (peek() ? std::string(1, peek()) : "end of file") + "'");
        take();
    }
    void want(const char *word) {
        if (!at(word)) fail(std::string("'") + word + "' expected");
        for (size_t i = strlen(word); i > 0; i--) take();
    }
    bool at_comment() const { return peek() == '/' && pos + 1 < src.size() && src[pos + 1] == '/'; }
    void rest_of_line() {
        while (peek() && peek() != '\n') take();
    }
    // whitespace (newlines included) and // comments
    void skip() {
        for (;;) {
            while (isspace((unsigned char)peek())) take();
            if (!at_comment()) return;
            rest_of_line();
        }
    }
    void want_space() {
        if (!isspace((unsigned char)peek())) fail("space expected");
        take();
    }
    // a statement ends at the end of its line; a trailing comment is allowed (end of file is accepted too)
    void end_of_statement() {
        while (peek() == ' ' || peek() == '\t' || peek() == '\r') take();
        if (at_comment()) rest_of_line();
        if (peek() == '\n' || peek() == '\0') {
            skip();
            return;
        }
        fail("newline expected");
    }
};

bool ident_start(char c) { return isalpha((unsigned char)c) || c == '_'; }

std::string identifier(Cursor &c) {
    if (!ident_start(c.peek())) c.fail(std::string("illegal character in identifier: '") + c.peek() + "'");
    std::string s;
    while (isalnum((unsigned char)c.peek()) || c.peek() == '_') s += c.take();
    return s;
}

std::string quoted(Cursor &c) {
    c.want('"');
    std::string s;
    while (c.peek() && c.peek() != '"') {
        char ch = c.take();
        if (ch == '\\') {
            const char e = c.take();
            if (e == '\\' || e == '"')
                ch = e;
            else if (e == 'n')
                ch = '\n';
            else
                c.fail("illegal escape sequence");
        }
        s += ch;
    }
    c.want('"');
    return s;
}

// core/parser.cpp:338-363: integer part and fraction digits are accumulated as integers and combined as
// n + frac / 10^k in double; no exponent syntax. The same evaluation keeps every literal's double (and hence float)
// value identical to what the reference reads from the same file.
double number(Cursor &c) {
    if (c.peek() == '-') {
        c.take();
        return -number(c);
    }
    if (!isdigit((unsigned char)c.peek())) c.fail("digit expected");
    long long n = 0, frac = 0;
    double p = 1.0;
    while (isdigit((unsigned char)c.peek())) n = n * 10 + (c.take() - '0');
    if (c.peek() == '.') {
        c.take();
        while (isdigit((unsigned char)c.peek())) {
            frac = frac * 10 + (c.take() - '0');
            p *= 10;
        }
    }
    return (double)n + (double)frac / p;
}

Value value(Cursor &c);

Value reference(Cursor &c) {
    c.want('$');
    std::vector<std::string> path{identifier(c)};
    while (c.peek() == '.') {
        c.take();
        path.push_back(identifier(c));
    }
    Module *m = c.mod;
    for (size_t i = 0; i + 1 < path.size(); i++) {
        auto it = m->submodules.find(path[i]);
        if (it == m->submodules.end()) c.fail("module " + m->name + " has no submodule named " + path[i]);
        m = it->second.get();
    }
    const std::string &leaf = path.back();
    if (path.size() == 1) {
        auto l = c.mod->locals.find(leaf);
        if (l != c.mod->locals.end()) return l->second;
    }
    auto e = m->exports.find(leaf);
    if (e == m->exports.end()) c.fail("module " + m->name + " has no exported variable named " + leaf);
    return e->second;
}

Value array(Cursor &c) {
    c.want('[');
    c.skip();
    Value v;
    v.kind = Value::Array;
    while (c.peek() && c.peek() != ']') {
        v.arr.push_back(value(c));
        c.skip();
        if (c.peek() != ']') c.want(',');
        c.skip();
    }
    c.want(']');
    return v;
}

Value object(Cursor &c) {
    auto o = std::make_shared<Object>();
    o->type = identifier(c);
    o->file = c.file;
    if (c.known && !c.known(o->type)) c.fail("unknown node type " + o->type);
    c.skip();
    c.want('{');
    c.skip();
    while (c.peek() && c.peek() != '}') {
        const std::string name = identifier(c);
        if (o->find(name)) c.fail("field " + name + " redefined");
        c.skip();
        c.want(':');
        c.skip();
        Value v = value(c);
        o->fields.emplace_back(name, std::move(v));
        c.skip();
        if (c.peek() != '}') c.want(',');
        c.skip();
    }
    c.want('}');
    Value v;
    v.kind = Value::Obj;
    v.obj = std::move(o);
    return v;
}

Value value(Cursor &c) {
    const char ch = c.peek();
    if (ch == '$') return reference(c);
    if (ch == '[') return array(c);
    if (ch == '"') {
        Value v;
        v.kind = Value::String;
        v.str = quoted(c);
        return v;
    }
    if (ch == '-' || isdigit((unsigned char)ch)) {
        Value v;
        v.kind = Value::Number;
        v.num = number(c);
        return v;
    }
    if (ident_start(ch)) {
        // the reference tests the prefix only (core/parser.cpp:253-259), so `trueX` reads as true followed by garbage
        if (c.at("true") || c.at("false")) {
            Value v;
            v.kind = Value::Bool;
            v.b = c.at("true");
            c.want(v.b ? "true" : "false");
            return v;
        }
        return object(c);
    }
    c.fail(std::string("stray token ") + (ch ? std::string(1, ch) : "end of file"));
}

void binding(Cursor &c, const char *keyword, std::map<std::string, Value> &table) {
    c.want(keyword);
    c.want_space();
    const std::string name = identifier(c);
    if (table.count(name)) c.fail(name + " is already defined");
    c.skip();
    c.want('=');
    c.skip();
    table[name] = value(c);
    c.end_of_statement();
}

void import(Cursor &c) {
    c.want("import");
    c.want_space();
    const std::string rel = quoted(c);
    const std::string path = join_path(dir_of(c.file), rel);
    {
        std::ifstream probe(path);
        if (!probe.good()) c.fail("module \"" + rel + "\" not found");
    }
    c.want_space();
    c.want("as");
    c.want_space();
    c.skip();
    const std::string alias = identifier(c);
    if (c.mod->submodules.count(alias)) c.fail(alias + " is already defined");
    c.end_of_statement();
    c.mod->submodules[alias] = parse_file(path, alias, c.known);
}

} // namespace

std::shared_ptr<Module> parse_string(const std::string &src, const std::string &file, const std::string &module_name,
                                     bool (*is_known_type)(const std::string &)) {
    auto mod = std::make_shared<Module>();
    mod->name = module_name, mod->file = file;
    Cursor c{src, file};
    c.mod = mod.get(), c.known = is_known_type;
    c.skip();
    while (c.peek()) {
        if (c.at("import"))
            import(c);
        else if (c.at("let"))
            binding(c, "let", mod->locals);
        else if (c.at("export"))
            binding(c, "export", mod->exports);
        else
            c.fail(std::string("stray token ") + c.peek());
    }
    return mod;
}

std::shared_ptr<Module> parse_file(const std::string &path, const std::string &module_name,
                                   bool (*is_known_type)(const std::string &)) {
    std::ifstream in(path, std::ios::binary);
    if (!in.good()) throw ParseError("cannot open " + path);
    const std::string src((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
    return parse_string(src, path, module_name, is_known_type);
}

} // namespace akb_host
