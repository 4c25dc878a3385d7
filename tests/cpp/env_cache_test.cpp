// Unit test of src/util/env.h (compiled and run by tests/test_package_api.py): a cached lookup keeps its value until the
// generation is bumped, every call site has its own slot, and unset / empty variables fall back to the default.
#include <cstdio>
#include <cstdlib>

#include "util/env.h"

static int read_a(int dflt) { return B2R_ENV_INT("B2R_TEST_A", dflt); }
static int read_a_elsewhere(int dflt) { return B2R_ENV_INT("B2R_TEST_A", dflt); }
static int read_b(int dflt) { return B2R_ENV_INT("B2R_TEST_B", dflt); }

#define CHECK(cond)                                             \
  do {                                                          \
    if (!(cond)) {                                              \
      fprintf(stderr, "FAILED line %d: %s\n", __LINE__, #cond); \
      return 1;                                                 \
    }                                                           \
  } while (0)

int main() {
  unsetenv("B2R_TEST_A");
  unsetenv("B2R_TEST_B");
  CHECK(read_a(7) == 7);          // unset -> default
  CHECK(read_a(9) == 9);          // the default is not cached, only the variable
  setenv("B2R_TEST_A", "42", 1);
  CHECK(read_a(7) == 7);          // still the cached "unset"
  CHECK(read_a_elsewhere(7) == 42);  // another call site reads on its first use
  b2r::env_refresh();
  CHECK(read_a(7) == 42);
  CHECK(read_a_elsewhere(7) == 42);  // every slot reads again on its first use after a refresh
  setenv("B2R_TEST_A", "43", 1);
  CHECK(read_a(7) == 42 && read_a_elsewhere(7) == 42);
  b2r::env_refresh();
  CHECK(read_a(7) == 43 && read_a_elsewhere(7) == 43);
  setenv("B2R_TEST_A", "", 1);    // empty counts as unset
  setenv("B2R_TEST_B", "-5", 1);
  b2r::env_refresh();
  CHECK(read_a(3) == 3);
  CHECK(read_b(0) == -5);
  unsetenv("B2R_TEST_B");
  CHECK(read_b(0) == -5);
  b2r::env_refresh();
  CHECK(read_b(11) == 11);
  printf("OK\n");
  return 0;
}
