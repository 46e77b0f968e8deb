/* oracle/refshim/boost/program_options.hpp -- TEST INFRASTRUCTURE: the subset of boost::program_options that the
 * reference's main uses (AutoBubStart3.cpp:161-196): options_description / add_options() with value<T>(&x),
 * ->default_value(), bool_switch(&b); command_line_parser(argc, argv).options(d).run(); store; notify;
 * variables_map::count.  "--long value", "--long=value", "-s value", "-svalue".  Boost is not in this image. */
#ifndef ABUB_REFSHIM_BOOST_PROGRAM_OPTIONS_HPP
#define ABUB_REFSHIM_BOOST_PROGRAM_OPTIONS_HPP

#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace boost {
namespace program_options {

struct error : public std::logic_error {
    explicit error(const std::string& w) : std::logic_error(w) {}
};

struct value_semantic {
    virtual ~value_semantic() {}
    virtual bool takes_arg() const = 0;
    virtual void apply(const std::string& text) = 0;
    virtual void apply_default() = 0;
};

template <typename T> struct typed_value : public value_semantic {
    T* dst;
    bool is_switch = false, has_default = false;
    T def{};
    explicit typed_value(T* d) : dst(d) {}
    typed_value* default_value(const T& v) { def = v; has_default = true; return this; }
    bool takes_arg() const override { return !is_switch; }
    void apply(const std::string& text) override {
        if (is_switch) { assign_switch(); return; }
        std::istringstream ss(text);
        T v{};
        read(ss, v, text);
        if (dst) *dst = v;
    }
    void apply_default() override { if (has_default && dst) *dst = def; }

private:
    template <typename U> static void read(std::istringstream& ss, U& v, const std::string& text) {
        if (!(ss >> v)) throw error("the argument ('" + text + "') for an option is invalid");
    }
    static void read(std::istringstream&, std::string& v, const std::string& text) { v = text; }
    void assign_switch() { assign_true(dst); }
    static void assign_true(bool* p) { if (p) *p = true; }
    template <typename U> static void assign_true(U*) {}
};

template <typename T> typed_value<T>* value(T* v) { return new typed_value<T>(v); }
inline typed_value<bool>* bool_switch(bool* v) {
    typed_value<bool>* t = new typed_value<bool>(v);
    t->is_switch = true;
    t->default_value(false);
    return t;
}

struct option_description {
    std::string long_name, short_name;
    std::shared_ptr<value_semantic> semantic;   /* null: a flag without storage ("help,h") */
};

class options_description;
class options_description_easy_init {
public:
    explicit options_description_easy_init(options_description* o) : owner(o) {}
    options_description_easy_init& operator()(const char* name, const char* description);
    options_description_easy_init& operator()(const char* name, value_semantic* s, const char* description);

private:
    options_description* owner;
};

class options_description {
public:
    explicit options_description(const std::string& caption = "") : caption_(caption) {}
    options_description_easy_init add_options() { return options_description_easy_init(this); }
    std::vector<option_description> options;
    const option_description* find(const std::string& name, bool is_short) const {
        for (const option_description& o : options)
            if ((is_short ? o.short_name : o.long_name) == name) return &o;
        return nullptr;
    }

private:
    std::string caption_;
};

inline void split_name(const char* name, option_description& o) {
    std::string n(name);
    const size_t comma = n.find(',');
    o.long_name = n.substr(0, comma);
    if (comma != std::string::npos) o.short_name = n.substr(comma + 1);
}
inline options_description_easy_init& options_description_easy_init::operator()(const char* name, const char*) {
    option_description o;
    split_name(name, o);
    owner->options.push_back(o);
    return *this;
}
inline options_description_easy_init& options_description_easy_init::operator()(const char* name, value_semantic* s, const char*) {
    option_description o;
    split_name(name, o);
    o.semantic.reset(s);
    owner->options.push_back(o);
    return *this;
}

struct parsed_options {
    const options_description* desc = nullptr;
    std::vector<std::pair<const option_description*, std::string> > found;
};

class command_line_parser {
public:
    command_line_parser(int argc, char** argv) { for (int i = 1; i < argc; i++) args.push_back(argv[i]); }
    command_line_parser& options(const options_description& d) { desc = &d; return *this; }
    parsed_options run() {
        parsed_options out;
        out.desc = desc;
        for (size_t i = 0; i < args.size(); i++) {
            const std::string& a = args[i];
            const option_description* o = nullptr;
            std::string val;
            bool have_val = false;
            if (a.rfind("--", 0) == 0) {
                std::string name = a.substr(2);
                const size_t eq = name.find('=');
                if (eq != std::string::npos) { val = name.substr(eq + 1); name = name.substr(0, eq); have_val = true; }
                o = desc->find(name, false);
                if (!o) throw error("unrecognised option '" + a + "'");
            } else if (a.size() >= 2 && a[0] == '-') {
                o = desc->find(a.substr(1, 1), true);
                if (!o) throw error("unrecognised option '" + a + "'");
                if (a.size() > 2) { val = a.substr(2); have_val = true; }
            } else {
                throw error("too many positional options have been specified on the command line");
            }
            const bool wants = o->semantic && o->semantic->takes_arg();
            if (wants && !have_val) {
                if (i + 1 >= args.size()) throw error("the required argument for option '" + a + "' is missing");
                val = args[++i];
            }
            out.found.push_back(std::make_pair(o, val));
        }
        return out;
    }

private:
    std::vector<std::string> args;
    const options_description* desc = nullptr;
};

class variables_map {
public:
    size_t count(const std::string& name) const { return seen.count(name); }
    std::map<std::string, std::string> seen;
    const options_description* desc = nullptr;
};

inline void store(const parsed_options& p, variables_map& vm) {
    vm.desc = p.desc;
    for (const option_description& o : p.desc->options)
        if (o.semantic) o.semantic->apply_default();
    for (const auto& f : p.found) {
        vm.seen[f.first->long_name] = f.second;
        if (f.first->semantic) f.first->semantic->apply(f.second);
    }
}
inline void notify(variables_map&) {}

}  // namespace program_options
}  // namespace boost
#endif
