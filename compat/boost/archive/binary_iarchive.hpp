#include "../serialization/serialization.hpp"
