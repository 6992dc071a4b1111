// TEST INFRASTRUCTURE — empty stand-in, see boost/log/trivial.hpp in this shim (reference: src/Logging.h:6-11)
#pragma once
