// prints a parsed YAML file as flattened "path = value" lines (test helper for hypre-mini-app_b200/app/yaml_lite.h)
#include <iostream>
#include "yaml_lite.h"
static void dump(const YAML::Node &n, const std::string &path) {
  if (n.IsMap()) {
    for (const std::string &k : n.keys()) dump(n[k], path.empty() ? k : path + "." + k);
  } else if (n.IsSequence()) {
    for (size_t i = 0; i < n.size(); ++i) dump(n[i], path + "[" + std::to_string(i) + "]");
  } else if (n.IsScalar()) {
    std::cout << path << " = " << n.as<std::string>() << "\n";
  } else {
    std::cout << path << " = ~\n";
  }
}
int main(int argc, char **argv) {
  try {
    YAML::Node root = YAML::LoadFile(argv[1]);
    dump(root, "");
    if (argc > 2) {  // typed access checks
      YAML::Node s = root["solver_settings"];
      std::cout << "tol*2 = " << 2 * s["tolerance"].as<double>() << "\n";
      std::cout << "kspace+1 = " << s["kspace"].as<int>() + 1 << "\n";
      std::cout << "missing = " << (s["nope"] ? 1 : 0) << "\n";
      std::cout << "flag = " << (root["linear_system"]["write_outputs"].as<bool>() ? 1 : 0) << "\n";
    }
  } catch (const std::exception &e) {
    std::cerr << "error: " << e.what() << "\n";
    return 1;
  }
  return 0;
}
