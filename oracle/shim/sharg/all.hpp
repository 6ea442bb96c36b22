// oracle/shim/sharg/all.hpp -- TEST INFRASTRUCTURE ONLY: the slice of sharg-parser (un-vendored dependency of the
// reference, bucket_map/CMakeLists.txt:75-80) that bucket_map/main.cpp:41-156 uses.
#pragma once
#include <cstdint>
#include <filesystem>
#include <functional>
#include <sstream>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

namespace sharg {

class parser_error : public std::runtime_error {
public:
    using std::runtime_error::runtime_error;
};

namespace detail {
struct default_validator {
    template <class T>
    void operator()(T const&) const {}
};
}   // namespace detail

class input_file_validator {
    std::vector<std::string> ext;

public:
    input_file_validator() = default;
    input_file_validator(std::vector<std::string> e) : ext(std::move(e)) {}
    void operator()(std::filesystem::path const& p) const {
        if (!std::filesystem::exists(p)) throw parser_error("The file " + p.string() + " does not exist!");
        if (ext.empty()) return;
        std::string e = p.extension().string();
        if (!e.empty()) e = e.substr(1);
        for (auto const& x : ext) if (x == e) return;
        throw parser_error("Expected one of the following valid extensions for " + p.string());
    }
    template <class T>
    void operator()(T const&) const {}
};

class output_file_validator {
    std::vector<std::string> ext;

public:
    output_file_validator() = default;
    output_file_validator(char const* e) : ext{std::string(e)} {}
    output_file_validator(std::vector<std::string> e) : ext(std::move(e)) {}
    void operator()(std::filesystem::path const& p) const {
        if (std::filesystem::exists(p)) throw parser_error("The file " + p.string() + " already exists!");
        if (ext.empty()) return;
        std::string e = p.extension().string();
        if (!e.empty()) e = e.substr(1);
        for (auto const& x : ext) if (x == e) return;
        throw parser_error("Expected one of the following valid extensions for " + p.string());
    }
    template <class T>
    void operator()(T const&) const {}
};

template <class validator_t = detail::default_validator>
struct config {
    char short_id{'\0'};
    std::string long_id{};
    std::string description{};
    std::string default_message{};
    bool advanced{false};
    bool hidden{false};
    bool required{false};
    validator_t validator{};
};

struct parser_meta_data {
    std::string app_name, version, short_description, author, email, date, url, short_copyright, long_copyright, citation;
    std::vector<std::string> description, synopsis, examples;
};

class parser {
    struct opt {
        char s; std::string l; bool flag; bool required; bool seen = false;
        std::function<void(std::string const&)> set;
    };
    std::vector<opt> opts;
    std::vector<std::string> argv_;

    template <class T>
    static void assign(T& v, std::string const& s) {
        if constexpr (std::is_same_v<T, std::string>) v = s;
        else if constexpr (std::is_same_v<T, std::filesystem::path>) v = s;
        else if constexpr (std::is_same_v<T, uint8_t> || std::is_same_v<T, int8_t>) {
            int x; std::istringstream is(s);
            if (!(is >> x) || !is.eof()) throw parser_error("Value parse failed for '" + s + "'");
            v = (T)x;
        } else {
            std::istringstream is(s);
            if (!(is >> v) || !is.eof()) throw parser_error("Value parse failed for '" + s + "'");
        }
    }

public:
    parser_meta_data info;
    parser(std::string const& name, int argc, char** argv) {
        info.app_name = name;
        for (int i = 1; i < argc; i++) argv_.push_back(argv[i]);
    }
    template <class cfg_t>
    void add_flag(bool& v, cfg_t const& c) {
        opts.push_back({c.short_id, c.long_id, true, false, false, [&v](std::string const&) { v = true; }});
    }
    template <class T, class cfg_t>
    void add_option(T& v, cfg_t const& c) {
        auto val = c.validator;
        opts.push_back({c.short_id, c.long_id, false, c.required, false, [&v, val](std::string const& s) {
                            assign(v, s);
                            val(v);
                        }});
    }
    void parse() {
        for (std::size_t i = 0; i < argv_.size(); i++) {
            std::string a = argv_[i], value;
            bool has_value = false;
            opt* o = nullptr;
            if (a.rfind("--", 0) == 0) {
                std::string name = a.substr(2);
                auto eq = name.find('=');
                if (eq != std::string::npos) { value = name.substr(eq + 1); name = name.substr(0, eq); has_value = true; }
                if (name == "version-check") { if (!has_value) i++; continue; }      // accepted and ignored
                for (auto& x : opts) if (x.l == name) o = &x;
            } else if (a.size() >= 2 && a[0] == '-') {
                for (auto& x : opts) if (x.s == a[1]) o = &x;
                if (a.size() > 2) { value = a.substr(2); has_value = true; }
            }
            if (!o) throw parser_error("Unknown option " + a + ". In case this is meant to be a non-option/argument/parameter, please specify the start of arguments with '--'.");
            if (o->flag) { o->set(""); o->seen = true; continue; }
            if (!has_value) {
                if (i + 1 >= argv_.size()) throw parser_error("Missing value for option " + a);
                value = argv_[++i];
            }
            o->set(value);
            o->seen = true;
        }
        for (auto const& x : opts)
            if (x.required && !x.seen) throw parser_error(std::string("Option -") + x.s + "/--" + x.l + " is required but not set.");
    }
};

}   // namespace sharg
