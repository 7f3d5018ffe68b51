// Process-wide bookkeeping of libpmdist: last-error text, launch counter, version.
#include <string>

#include "../include/pmdist.h"
#define PMD_API __attribute__((visibility("default")))

namespace pmd {
static thread_local std::string g_error;
unsigned long long g_launch_count = 0;
void set_error(const std::string& msg) { g_error = msg; }
const char* last_error_cstr() { return g_error.c_str(); }
}  // namespace pmd

extern "C" PMD_API const char* pmd_last_error(void) { return pmd::last_error_cstr(); }
extern "C" PMD_API unsigned long long pmd_launch_count(void) { return pmd::g_launch_count; }
extern "C" PMD_API int pmd_version(void) { return 100; }
