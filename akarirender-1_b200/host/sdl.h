// sdl.h — reader for the `.akari` scene description language.
//
// Replaces akari::sdl::Parser (core/parser.cpp:150-363) for the node vocabulary of the Path hot path. Grammar accepted
// (and rejected) as the reference does: a module is a sequence of newline-terminated statements
//     import "file.akari" as name        (path relative to the importing file, core/parser.cpp:167-193)
//     let id = value                      (module-local)
//     export id = value
// with `//` line comments; values are `$id` / `$mod.id` references (locals first, then exports; only a bare `$id`
// consults the locals, core/parser.cpp:224-247), `[v, ...]` arrays (trailing comma allowed), "strings" with the
// escapes \\ \n \", numbers `-?digits[.digits]` (no exponent), true / false, and objects `Type { field: value, ... }`.
// References resolve to the SAME object (shared), which is what lets one material node serve several meshes.
//
// This layer only builds the value tree; scene_file.cpp gives the node types their meaning.
#pragma once
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace akb_host {

struct ParseError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

struct Object;

struct Value {
    enum Kind { Null, Bool, Number, String, Array, Obj } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Value> arr;
    std::shared_ptr<Object> obj;

    bool is_number() const { return kind == Number; }
    bool is_array() const { return kind == Array; }
    bool is_string() const { return kind == String; }
    bool is_object() const { return kind == Obj; }
};

// One `Type { ... }` node. Fields keep file order; a field may not be given twice (core/parser.cpp:276-279).
struct Object {
    std::string type;
    std::string file; // the file that defined it: relative paths inside resolve against its directory
    std::vector<std::pair<std::string, Value>> fields;
    const Value *find(const std::string &name) const {
        for (auto &f : fields)
            if (f.first == name) return &f.second;
        return nullptr;
    }
};

struct Module {
    std::string name, file;
    std::map<std::string, Value> exports, locals;
    std::map<std::string, std::shared_ptr<Module>> submodules;
};

// `is_known_type` is asked for every object's type name while parsing (the reference creates the node at that point
// and fails on an unregistered name, core/nodes/scenegraph.cpp:45-60).
std::shared_ptr<Module> parse_file(const std::string &path, const std::string &module_name,
                                   bool (*is_known_type)(const std::string &));
std::shared_ptr<Module> parse_string(const std::string &src, const std::string &file, const std::string &module_name,
                                     bool (*is_known_type)(const std::string &));

std::string dir_of(const std::string &path);
std::string join_path(const std::string &dir, const std::string &rel);

} // namespace akb_host
