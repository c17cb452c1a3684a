// TEST INFRASTRUCTURE: stand-in for the sliver of boost::format the reference uses
// (sequential %s / %d directives, operator%, str(), operator<<).
#pragma once
#include <string>
#include <sstream>
#include <vector>
#include <algorithm>
#include <iostream>
namespace boost {
class format {
public:
    format(const char* f) : fmt_(f) {}
    format(const std::string& f) : fmt_(f) {}
    template <class T> format& operator%(const T& v) {
        std::ostringstream os; os << v; args_.push_back(os.str()); return *this;
    }
    std::string str() const {
        std::string out; size_t k = 0;
        for (size_t i = 0; i < fmt_.size(); i++) {
            if (fmt_[i] == '%' && i + 1 < fmt_.size()) {
                char c = fmt_[i + 1];
                if (c == '%') { out += '%'; i++; continue; }
                if (c == 's' || c == 'd' || c == 'u' || c == 'f') {
                    if (k < args_.size()) out += args_[k];
                    k++; i++; continue;
                }
            }
            out += fmt_[i];
        }
        return out;
    }
private:
    std::string fmt_;
    std::vector<std::string> args_;
};
inline std::ostream& operator<<(std::ostream& os, const format& f) { return os << f.str(); }
}
