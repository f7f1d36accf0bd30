#pragma once
#include <filesystem>
#include <chrono>
#include <ctime>
#include <string>
#include "system/error_code.hpp"
namespace boost { namespace filesystem {
// include/mitsuba/core/fwd.h forward-declares `class path`, so it must be a real class
class path : public std::filesystem::path {
public:
    path() {}
    path(const std::filesystem::path &p) : std::filesystem::path(p) {}
    path(const std::string &s) : std::filesystem::path(s) {}
    path(const char *s) : std::filesystem::path(s) {}
    path filename() const { return std::filesystem::path::filename(); }
    path parent_path() const { return std::filesystem::path::parent_path(); }
    path extension() const { return std::filesystem::path::extension(); }
    path stem() const { return std::filesystem::path::stem(); }
    path &replace_extension(const path &e = path()) { std::filesystem::path::replace_extension(e); return *this; }
    path operator/(const path &o) const { return path(static_cast<const std::filesystem::path &>(*this) / static_cast<const std::filesystem::path &>(o)); }
    path &operator/=(const path &o) { std::filesystem::path::operator/=(o); return *this; }
    bool is_complete() const { return is_absolute(); }
    std::string file_string() const { return string(); }
    path leaf() const { return filename(); }
    path branch_path() const { return parent_path(); }
};
inline bool exists(const path &p) { std::error_code ec; return std::filesystem::exists(p, ec); }
inline bool is_directory(const path &p) { std::error_code ec; return std::filesystem::is_directory(p, ec); }
inline bool is_regular_file(const path &p) { std::error_code ec; return std::filesystem::is_regular_file(p, ec); }
inline path absolute(const path &p) { return path(std::filesystem::absolute(p)); }
inline path complete(const path &p) { return absolute(p); }
inline path current_path() { return path(std::filesystem::current_path()); }
inline bool create_directory(const path &p) { std::error_code ec; return std::filesystem::create_directory(p, ec); }
inline bool create_directories(const path &p) { std::error_code ec; return std::filesystem::create_directories(p, ec); }
inline bool remove(const path &p) { std::error_code ec; return std::filesystem::remove(p, ec); }
inline uintmax_t file_size(const path &p) { std::error_code ec; return std::filesystem::file_size(p, ec); }
inline void rename(const path &a, const path &b) { std::filesystem::rename(a, b); }
inline std::time_t last_write_time(const path &p) { return 0; }
inline path canonical(const path &p) { return path(std::filesystem::weakly_canonical(p)); }
inline void resize_file(const path &p, uintmax_t n) { std::filesystem::resize_file(p, n); }
inline bool exists(const path &p, std::error_code &ec) { return std::filesystem::exists(p, ec); }
inline std::time_t last_write_time(const path &p, std::error_code &ec) {
    auto t = std::filesystem::last_write_time(p, ec);
    if (ec) return 0;
    return (std::time_t)std::chrono::duration_cast<std::chrono::seconds>(t.time_since_epoch()).count();
}
using std::filesystem::directory_iterator;
} }
