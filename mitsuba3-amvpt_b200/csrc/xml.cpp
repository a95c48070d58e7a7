// xml.cpp -- placeholder, replaced below
#include "host.h"
namespace amvpt {
Scene *load_xml_file(const std::string &, const std::vector<std::pair<std::string, std::string>> &) { AM_THROW("xml loader not built yet"); }
Scene *load_xml_string(const std::string &, const std::string &, const std::vector<std::pair<std::string, std::string>> &) { AM_THROW("xml loader not built yet"); }
}
