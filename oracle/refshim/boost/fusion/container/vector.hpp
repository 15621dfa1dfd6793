#include "cnf2_refshim.hpp"
