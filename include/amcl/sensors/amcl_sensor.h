/* Forwarder: "amcl/sensors/amcl_sensor.h" resolves to the combined C++ ABI header. */
#include "../../amcl_b200/amcl_sensors.hpp"
