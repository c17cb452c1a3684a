// TEST INFRASTRUCTURE: stand-in for boost::initialized_value (value-initialises any type).
#pragma once
namespace boost {
struct initialized_value_t {
    template <class T> operator T() const { return T(); }
};
static const initialized_value_t initialized_value = initialized_value_t();
}
