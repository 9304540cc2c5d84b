/* Minimal boost::optional stand-in for the two uses in the reference's zklaim/snark.cpp
 * (return type of generate_proof, comparison with boost::none, .get()).
 * TEST INFRASTRUCTURE only (oracle build). */
#ifndef ZKB_ORACLE_BOOST_OPTIONAL_SHIM
#define ZKB_ORACLE_BOOST_OPTIONAL_SHIM
#include <utility>
namespace boost {
struct none_t {};
static const none_t none = none_t();
template <typename T> class optional {
    bool has_;
    T val_;
public:
    optional() : has_(false), val_() {}
    optional(none_t) : has_(false), val_() {}
    optional(const T &v) : has_(true), val_(v) {}
    optional(T &&v) : has_(true), val_(std::move(v)) {}
    bool operator==(none_t) const { return !has_; }
    bool operator!=(none_t) const { return has_; }
    explicit operator bool() const { return has_; }
    T &get() { return val_; }
    const T &get() const { return val_; }
};
}
#endif
