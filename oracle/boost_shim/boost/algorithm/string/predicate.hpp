#ifndef SMC_BOOST_SHIM_PREDICATE
#define SMC_BOOST_SHIM_PREDICATE
#include <string>
namespace boost { namespace algorithm {
inline bool ends_with(const std::string& s, const std::string& suf) {
  return s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0;
}
}}
#endif
