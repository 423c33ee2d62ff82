// TEST INFRASTRUCTURE ONLY -- a small working implementation of the slice of Boost.Program_options that the reference's option
// classes use (src/SpydrPick_options.cpp, src/ARACNE_options.cpp, externals/apegrunt/src/Apegrunt_options.cpp, src/SpydrPick.cpp:65-85):
// options_description / add_options() chains, value<T>(ptr)->default_value()->notifier()->composing(), bool_switch, variables_map,
// positional_options_description, command_line_parser, store, notify. Long options (--name value, --name=value, unambiguous
// prefixes), short options (-v, -t 4, -t4), "--" terminator, positional arguments. Errors throw std::exception subclasses whose
// what() reads like Boost's ("unrecognised option '--x'").
#pragma once
#include <cstddef>
#include <functional>
#include <map>
#include <memory>
#include <ostream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace boost { namespace program_options {

class error : public std::logic_error { public: explicit error(const std::string& w) : std::logic_error(w) {} };
class unknown_option : public error { public: explicit unknown_option(const std::string& n) : error("unrecognised option '" + n + "'") {} };
class invalid_option_value : public error { public: invalid_option_value(const std::string& n, const std::string& v) : error("the argument ('" + v + "') for option '--" + n + "' is invalid") {} };
class required_argument_missing : public error { public: explicit required_argument_missing(const std::string& n) : error("the required argument for option '--" + n + "' is missing") {} };
class ambiguous_option : public error { public: explicit ambiguous_option(const std::string& n) : error("option '--" + n + "' is ambiguous") {} };
class too_many_positional_options_error : public error { public: too_many_positional_options_error() : error("too many positional options have been specified on the command line") {} };

// type-erased value holder (boost::any stand-in)
class variable_value {
public:
    variable_value() = default;
    template <class T> explicit variable_value(const T& v, bool defaulted) : m_p(std::make_shared<holder<T>>(v)), m_defaulted(defaulted) {}
    bool empty() const { return !m_p; }
    bool defaulted() const { return m_defaulted; }
    template <class T> const T& as() const { auto* h = dynamic_cast<holder<T>*>(m_p.get()); if (!h) throw error("bad variable_value cast"); return h->v; }
    template <class T> T& as() { auto* h = dynamic_cast<holder<T>*>(m_p.get()); if (!h) throw error("bad variable_value cast"); return h->v; }
private:
    struct base { virtual ~base() = default; };
    template <class T> struct holder : base { explicit holder(const T& x) : v(x) {} T v; };
    std::shared_ptr<base> m_p; bool m_defaulted = false;
};

class value_semantic {
public:
    virtual ~value_semantic() = default;
    virtual bool takes_argument() const = 0;
    virtual bool is_composing() const = 0;
    virtual bool has_default() const = 0;
    virtual std::string default_text() const = 0;
    virtual variable_value make_default() const = 0;
    virtual void parse_into(variable_value& slot, const std::string& name, const std::vector<std::string>& tokens) const = 0;
    virtual void notify(const variable_value& v) const = 0;
};

namespace detail_shim {
template <class T> struct is_vector : std::false_type {};
template <class T, class A> struct is_vector<std::vector<T, A>> : std::true_type {};
template <class T> inline bool lexical(const std::string& s, T& out) { std::istringstream is(s); is >> out; return !is.fail() && (is.peek(), is.eof()); }
template <> inline bool lexical<std::string>(const std::string& s, std::string& out) { out = s; return true; }
template <class T> inline std::string to_text(const T& v) { std::ostringstream os; os << v; return os.str(); }
template <class T, class A> inline std::string to_text(const std::vector<T, A>&) { return std::string(); }
}

template <class T> class typed_value : public value_semantic {
public:
    explicit typed_value(T* store_to) : m_store_to(store_to) {}
    typed_value* default_value(const T& v) { m_default = std::make_shared<T>(v); m_default_text = detail_shim::to_text(v); return this; }
    typed_value* default_value(const T& v, const std::string& text) { m_default = std::make_shared<T>(v); m_default_text = text; return this; }
    typed_value* implicit_value(const T& v) { m_implicit = std::make_shared<T>(v); return this; }
    template <class F> typed_value* notifier(F f) { m_notifier = [f](const T& v) { f(v); }; return this; }
    typed_value* composing() { m_composing = true; return this; }
    typed_value* multitoken() { return this; }
    typed_value* zero_tokens() { m_zero_tokens = true; return this; }
    typed_value* required() { return this; }

    bool takes_argument() const override { return !m_zero_tokens; }
    bool is_composing() const override { return m_composing; }
    bool has_default() const override { return bool(m_default); }
    std::string default_text() const override { return m_default_text; }
    variable_value make_default() const override { return variable_value(*m_default, true); }
    void parse_into(variable_value& slot, const std::string& name, const std::vector<std::string>& tokens) const override { parse_impl(slot, name, tokens, detail_shim::is_vector<T>()); }
    void notify(const variable_value& v) const override {
        const T& x = v.as<T>();
        if (m_store_to) *m_store_to = x;
        if (m_notifier) m_notifier(x);
    }
private:
    void parse_impl(variable_value& slot, const std::string& name, const std::vector<std::string>& tokens, std::false_type) const {
        if (m_zero_tokens) { slot = variable_value(make_true(), false); return; }
        if (tokens.empty()) { if (m_implicit) { slot = variable_value(*m_implicit, false); return; } throw required_argument_missing(name); }
        T v{}; if (!detail_shim::lexical(tokens.back(), v)) throw invalid_option_value(name, tokens.back());
        slot = variable_value(v, false);
    }
    void parse_impl(variable_value& slot, const std::string& name, const std::vector<std::string>& tokens, std::true_type) const {
        if (tokens.empty()) throw required_argument_missing(name);
        T acc; if (!slot.empty() && !slot.defaulted() && m_composing) acc = slot.as<T>();
        for (const auto& t : tokens) { typename T::value_type e{}; if (!detail_shim::lexical(t, e)) throw invalid_option_value(name, t); acc.push_back(e); }
        slot = variable_value(acc, false);
    }
    static T make_true() { return make_true_impl(std::is_same<T, bool>()); }
    static T make_true_impl(std::true_type) { return true; }
    static T make_true_impl(std::false_type) { return T{}; }
    T* m_store_to; std::shared_ptr<T> m_default, m_implicit; std::string m_default_text;
    std::function<void(const T&)> m_notifier; bool m_composing = false, m_zero_tokens = false;
};

template <class T> typed_value<T>* value() { return new typed_value<T>(nullptr); }
template <class T> typed_value<T>* value(T* store_to) { return new typed_value<T>(store_to); }
inline typed_value<bool>* bool_switch(bool* store_to = nullptr) { auto* v = new typed_value<bool>(store_to); v->default_value(false); v->zero_tokens(); return v; }

struct option_description {
    std::string long_name, short_name, description;
    std::shared_ptr<const value_semantic> semantic;            // null: a plain flag ("help,h")
};

class options_description;
class options_description_easy_init {
public:
    explicit options_description_easy_init(options_description* owner) : m_owner(owner) {}
    options_description_easy_init& operator()(const char* name, const char* description);
    options_description_easy_init& operator()(const char* name, const value_semantic* s);
    options_description_easy_init& operator()(const char* name, const value_semantic* s, const char* description);
private:
    options_description* m_owner;
};

class options_description {
public:
    options_description() = default;
    explicit options_description(const std::string& caption) : m_caption(caption) {}
    options_description_easy_init add_options() { return options_description_easy_init(this); }
    options_description& add(const options_description& other) { for (const auto& o : other.m_options) m_options.push_back(o); return *this; }
    void add(const std::shared_ptr<option_description>& o) { m_options.push_back(o); }
    const std::vector<std::shared_ptr<option_description>>& options() const { return m_options; }
    const option_description* find_long(const std::string& name, bool allow_prefix = true) const {
        for (const auto& o : m_options) if (o->long_name == name) return o.get();
        if (!allow_prefix) return nullptr;
        const option_description* hit = nullptr;
        for (const auto& o : m_options) if (o->long_name.compare(0, name.size(), name) == 0) { if (hit) throw ambiguous_option(name); hit = o.get(); }
        return hit;
    }
    const option_description* find_short(char ch) const {
        for (const auto& o : m_options) if (o->short_name.size() == 1 && o->short_name[0] == ch) return o.get();
        return nullptr;
    }
    friend std::ostream& operator<<(std::ostream& os, const options_description& d) {
        if (!d.m_caption.empty()) os << d.m_caption << ":\n";
        for (const auto& o : d.m_options) {
            std::string head = "  ";
            if (!o->short_name.empty()) head += "-" + o->short_name + " [ --" + o->long_name + " ]"; else head += "--" + o->long_name;
            if (o->semantic && o->semantic->takes_argument()) { head += " arg"; if (o->semantic->has_default() && !o->semantic->default_text().empty()) head += " (=" + o->semantic->default_text() + ")"; }
            os << head; if (head.size() < 40) os << std::string(40 - head.size(), ' '); else os << "\n" << std::string(40, ' ');
            os << o->description << "\n";
        }
        return os;
    }
private:
    std::string m_caption; std::vector<std::shared_ptr<option_description>> m_options;
};

namespace detail_shim {
inline std::shared_ptr<option_description> make_option(const char* name, const value_semantic* s, const char* description) {
    auto o = std::make_shared<option_description>();
    const std::string n(name); const std::size_t comma = n.find(',');
    o->long_name = n.substr(0, comma); if (comma != std::string::npos) o->short_name = n.substr(comma + 1);
    o->description = description ? description : ""; o->semantic.reset(s);
    return o;
}
}
inline options_description_easy_init& options_description_easy_init::operator()(const char* name, const char* description) { m_owner->add(detail_shim::make_option(name, nullptr, description)); return *this; }
inline options_description_easy_init& options_description_easy_init::operator()(const char* name, const value_semantic* s) { m_owner->add(detail_shim::make_option(name, s, nullptr)); return *this; }
inline options_description_easy_init& options_description_easy_init::operator()(const char* name, const value_semantic* s, const char* description) { m_owner->add(detail_shim::make_option(name, s, description)); return *this; }

class positional_options_description {
public:
    positional_options_description& add(const char* name, int max_count) { m_names.emplace_back(name, max_count); return *this; }
    // name of the k-th positional argument; empty when there are too many
    std::string name_for_position(std::size_t k) const {
        std::size_t base = 0;
        for (const auto& e : m_names) { if (e.second < 0 || k < base + std::size_t(e.second)) return e.first; base += std::size_t(e.second); }
        return std::string();
    }
private:
    std::vector<std::pair<std::string, int>> m_names;
};

struct parsed_option { std::string name; std::vector<std::string> tokens; };
struct parsed_options { const options_description* description = nullptr; std::vector<parsed_option> options; };

class command_line_parser {
public:
    command_line_parser(int argc, const char* const* argv) { for (int k = 1; k < argc; ++k) m_args.emplace_back(argv[k]); }
    explicit command_line_parser(const std::vector<std::string>& args) : m_args(args) {}
    command_line_parser& options(const options_description& d) { m_desc = &d; return *this; }
    command_line_parser& positional(const positional_options_description& p) { m_pos = &p; return *this; }
    command_line_parser& allow_unregistered() { return *this; }
    parsed_options run() const {
        parsed_options out; out.description = m_desc; std::size_t npos = 0; bool only_positional = false;
        auto positional = [&](const std::string& tok) {
            const std::string nm = m_pos ? m_pos->name_for_position(npos) : std::string();
            if (nm.empty()) throw too_many_positional_options_error();
            ++npos; out.options.push_back({nm, {tok}});
        };
        for (std::size_t k = 0; k < m_args.size(); ++k) {
            const std::string& a = m_args[k];
            if (only_positional || a.size() < 2 || a[0] != '-') { positional(a); continue; }
            if (a == "--") { only_positional = true; continue; }
            if (a[1] == '-') {
                std::string name = a.substr(2), val; bool has_val = false;
                const std::size_t eq = name.find('=');
                if (eq != std::string::npos) { val = name.substr(eq + 1); name = name.substr(0, eq); has_val = true; }
                const option_description* o = m_desc ? m_desc->find_long(name) : nullptr;
                if (!o) throw unknown_option("--" + name);
                parsed_option po; po.name = o->long_name;
                if (o->semantic && o->semantic->takes_argument()) {
                    if (has_val) po.tokens.push_back(val);
                    else if (k + 1 < m_args.size()) po.tokens.push_back(m_args[++k]);
                    else throw required_argument_missing(o->long_name);
                } else if (has_val) throw error("option '--" + o->long_name + "' does not take any arguments");
                out.options.push_back(po);
            } else {
                // short options, possibly grouped ("-vt4")
                for (std::size_t c = 1; c < a.size(); ++c) {
                    const option_description* o = m_desc ? m_desc->find_short(a[c]) : nullptr;
                    if (!o) throw unknown_option(std::string("-") + a[c]);
                    parsed_option po; po.name = o->long_name;
                    if (o->semantic && o->semantic->takes_argument()) {
                        if (c + 1 < a.size()) po.tokens.push_back(a.substr(c + 1));
                        else if (k + 1 < m_args.size()) po.tokens.push_back(m_args[++k]);
                        else throw required_argument_missing(o->long_name);
                        out.options.push_back(po); break;
                    }
                    out.options.push_back(po);
                }
            }
        }
        return out;
    }
private:
    std::vector<std::string> m_args; const options_description* m_desc = nullptr; const positional_options_description* m_pos = nullptr;
};

class variables_map {
public:
    std::size_t count(const std::string& name) const { return m_values.count(name); }
    const variable_value& operator[](const std::string& name) const { static const variable_value none; auto it = m_values.find(name); return it == m_values.end() ? none : it->second; }
    bool empty() const { return m_values.empty(); }
    // used by store / notify
    std::map<std::string, variable_value> m_values;
    std::map<std::string, std::shared_ptr<const value_semantic>> m_semantics;
};

inline void store(const parsed_options& parsed, variables_map& vm) {
    const options_description* d = parsed.description;
    for (const auto& po : parsed.options) {
        const option_description* o = d ? d->find_long(po.name, false) : nullptr;
        if (!o) throw unknown_option("--" + po.name);
        variable_value& slot = vm.m_values[o->long_name];
        if (!o->semantic) { slot = variable_value(std::string(), false); continue; }
        if (!slot.empty() && !slot.defaulted() && !o->semantic->is_composing()) throw error("option '--" + o->long_name + "' cannot be specified more than once");
        o->semantic->parse_into(slot, o->long_name, po.tokens);
        vm.m_semantics[o->long_name] = o->semantic;
    }
    if (d) for (const auto& o : d->options()) {
        if (!o->semantic || !o->semantic->has_default() || vm.m_values.count(o->long_name)) continue;
        vm.m_values[o->long_name] = o->semantic->make_default();
        vm.m_semantics[o->long_name] = o->semantic;
    }
}

inline void notify(variables_map& vm) {
    for (const auto& kv : vm.m_semantics) { auto it = vm.m_values.find(kv.first); if (it != vm.m_values.end() && !it->second.empty()) kv.second->notify(it->second); }
}

}} // namespace boost::program_options
