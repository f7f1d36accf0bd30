#pragma once
#include <string>
#include <algorithm>
#include <cctype>
#include <vector>
namespace boost {
inline std::string to_lower_copy(std::string s) { std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char)std::tolower(c); }); return s; }
inline std::string to_upper_copy(std::string s) { std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char)std::toupper(c); }); return s; }
inline void to_lower(std::string &s) { s = to_lower_copy(s); }
inline void to_upper(std::string &s) { s = to_upper_copy(s); }
inline bool ends_with(const std::string &s, const std::string &e) { return s.size() >= e.size() && s.compare(s.size() - e.size(), e.size(), e) == 0; }
inline bool starts_with(const std::string &s, const std::string &e) { return s.size() >= e.size() && s.compare(0, e.size(), e) == 0; }
inline bool iequals(const std::string &a, const std::string &b) { return to_lower_copy(a) == to_lower_copy(b); }
inline std::string trim_copy(const std::string &s) { size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n"); return a == std::string::npos ? std::string() : s.substr(a, b - a + 1); }
inline void trim(std::string &s) { s = trim_copy(s); }
inline void replace_all(std::string &s, const std::string &from, const std::string &to) { if (from.empty()) return; size_t p = 0; while ((p = s.find(from, p)) != std::string::npos) { s.replace(p, from.size(), to); p += to.size(); } }
inline bool contains(const std::string &s, const std::string &t) { return s.find(t) != std::string::npos; }
namespace algorithm { using boost::to_lower_copy; using boost::to_upper_copy; using boost::ends_with; using boost::starts_with; using boost::trim_copy; }
}
