/* Stand-in for MVE's <util/exception.h>: util::FileException(filename, message).  TEST INFRASTRUCTURE ONLY. */
#ifndef ORACLE_REF_SHIM_UTIL_EXCEPTION_H
#define ORACLE_REF_SHIM_UTIL_EXCEPTION_H
#include <stdexcept>
#include <string>
namespace util {
class FileException : public std::runtime_error {
public:
    FileException(std::string const & filename, std::string const & msg) : std::runtime_error(filename + ": " + msg) {}
};
}
#endif
