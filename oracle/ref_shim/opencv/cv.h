// OpenCV header stand-in (oracle/ref_shim/cvshim.hpp): test infrastructure for compiling the unmodified reference.
#include "../cvshim.hpp"
