/* oracle/refshim/boost/filesystem.hpp -- TEST INFRASTRUCTURE: the subset of boost::filesystem (v3 semantics) that the
 * reference's ParseFolder/RawParser.cpp uses; directory access goes through std::filesystem.  Boost is not in this
 * image.  Note operator/=: Boost.Filesystem v3 APPENDS an operand that begins with a separator (the reference appends
 * ImageFolder = "/Images/", RawParser.cpp:36,143), where std::filesystem would replace the whole path. */
#ifndef ABUB_REFSHIM_BOOST_FILESYSTEM_HPP
#define ABUB_REFSHIM_BOOST_FILESYSTEM_HPP
#include <filesystem>
#include <ostream>
#include <string>
namespace boost {
namespace filesystem {

class path {
public:
    path() {}
    path(const std::string& s) : s_(s) {}
    path(const char* s) : s_(s) {}
    path& operator/=(const path& p) {
        if (p.s_.empty()) return *this;
        if (p.s_[0] != '/' && !s_.empty() && s_.back() != '/') s_ += '/';
        s_ += p.s_;
        return *this;
    }
    path& operator+=(const std::string& x) { s_ += x; return *this; }
    const std::string& native() const { return s_; }
    const std::string& string() const { return s_; }
    const char* c_str() const { return s_.c_str(); }
    path filename() const { return path(std::filesystem::path(s_).filename().string()); }
    path parent_path() const { return path(std::filesystem::path(s_).parent_path().string()); }
    bool empty() const { return s_.empty(); }

private:
    std::string s_;
};
inline path operator/(const path& a, const path& b) { path r(a); r /= b; return r; }
inline std::ostream& operator<<(std::ostream& o, const path& p) { return o << '"' << p.native() << '"'; }

class directory_entry {
public:
    explicit directory_entry(const std::string& s) : p_(s) {}
    const filesystem::path& path() const { return p_; }

private:
    filesystem::path p_;
};

/* range-for support: `for (const directory_entry& x : directory_iterator(dir))` */
class directory_iterator {
public:
    directory_iterator() {}
    explicit directory_iterator(const path& p) : it_(std::filesystem::path(p.native())) { load(); }
    bool operator!=(const directory_iterator& o) const { return it_ != o.it_; }
    directory_iterator& operator++() { ++it_; load(); return *this; }
    const directory_entry& operator*() const { return cur_; }

private:
    void load() { if (it_ != std::filesystem::directory_iterator()) cur_ = directory_entry(it_->path().string()); }
    std::filesystem::directory_iterator it_;
    directory_entry cur_{""};
};
inline directory_iterator begin(directory_iterator it) { return it; }
inline directory_iterator end(const directory_iterator&) { return directory_iterator(); }

inline bool exists(const path& p) { std::error_code ec; return std::filesystem::exists(std::filesystem::path(p.native()), ec); }
inline bool is_directory(const path& p) { std::error_code ec; return std::filesystem::is_directory(std::filesystem::path(p.native()), ec); }

}  // namespace filesystem
}  // namespace boost
#endif
