// TEST INFRASTRUCTURE ONLY (oracle build).  Stand-in for ut-amrl/config_reader
// (config_reader.h:73-177, macros.h:39-60) so that the UNMODIFIED kinematic
// model sources compile without Lua: CONFIG_<TYPE>(name, key) still declares
// `const T& CONFIG_name`, bound to a process-wide slot per key that the bridge
// (oracle/ref_bridge.cpp) fills instead of a Lua file.  ConfigReader itself
// reads nothing.
#ifndef UMRS_ORACLE_CONFIG_READER_SHIM
#define UMRS_ORACLE_CONFIG_READER_SHIM
#include <map>
#include <string>
#include <vector>

namespace config_reader {

template <typename T>
T& Slot(const std::string& key) {
  static std::map<std::string, T> slots;   // node addresses are stable
  return slots[key];
}

class ConfigReader {
 public:
  ConfigReader() = delete;
  explicit ConfigReader(const std::vector<std::string>&) {}
};

}  // namespace config_reader

#define CONFIG_INT(name, key) const int& CONFIG_##name = ::config_reader::Slot<int>(key)
#define CONFIG_UINT(name, key) const unsigned int& CONFIG_##name = ::config_reader::Slot<unsigned int>(key)
#define CONFIG_DOUBLE(name, key) const double& CONFIG_##name = ::config_reader::Slot<double>(key)
#define CONFIG_FLOAT(name, key) const float& CONFIG_##name = ::config_reader::Slot<float>(key)
#define CONFIG_STRING(name, key) const std::string& CONFIG_##name = ::config_reader::Slot<std::string>(key)
#define CONFIG_BOOL(name, key) const bool& CONFIG_##name = ::config_reader::Slot<bool>(key)
#endif
