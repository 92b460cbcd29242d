// oracle/stubs/pfs/parser_error.hpp -- TEST INFRASTRUCTURE ONLY (see procfs.hpp): MonData.h:18
// includes it; nothing the harness instantiates uses it.
#pragma once
