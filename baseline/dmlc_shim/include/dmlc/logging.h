/*! Minimal dmlc/logging.h stand-in for the baseline build: only dmlc::Error is needed. */
#ifndef DMLC_LOGGING_H_
#define DMLC_LOGGING_H_
#include <stdexcept>
#include <string>
#include "./base.h"
namespace dmlc {
struct Error : public std::runtime_error {
  explicit Error(const std::string &s) : std::runtime_error(s) {}
};
}  // namespace dmlc
#endif  // DMLC_LOGGING_H_
