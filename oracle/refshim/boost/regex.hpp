/* oracle/refshim/boost/regex.hpp -- TEST INFRASTRUCTURE: boost::regex / boost::regex_match as used by the reference's
 * ParseFolder/RawParser.cpp:145-150 ("^.*cam<N>.*(png|bmp)", whole-string match), mapped onto std::regex. */
#ifndef ABUB_REFSHIM_BOOST_REGEX_HPP
#define ABUB_REFSHIM_BOOST_REGEX_HPP
#include <regex>
#include <string>
namespace boost {
typedef std::regex regex;
inline bool regex_match(const std::string& s, const regex& re) { return std::regex_match(s, re); }
}  // namespace boost
#endif
