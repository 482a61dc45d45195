// global_config.json of the reference (config.{hpp,cc}, global_config.json.bak), parsed ONCE and
// cached in memory.  The reference re-parses ./global_config.json on every get_config() call and
// passes DENSE_MATRIX_SIZE / VECTOR_WIDTH between stages by rewriting the file (config.cc:17-40,
// token_test.cc:1630,1046); here set_config() updates the cache and, only when a file was loaded
// and write-back is on, the file too.  Keys keep the reference's names and meaning.
#pragma once
#include <map>
#include <string>

namespace gs {

class config_store {
  public:
    config_store(); // defaults = global_config.json.bak with HALF=false
    // load ./global_config.json style file (flat JSON object of strings, numbers, booleans)
    void load(const std::string &path);
    bool has(const std::string &key) const { return kv_.count(key) != 0; }
    std::string as_string(const std::string &key) const;
    double as_float(const std::string &key) const;
    long as_integer(const std::string &key) const { return (long)as_float(key); }
    bool as_bool(const std::string &key) const; // a missing key reads as false, like configor's null value
    void set(const std::string &key, const std::string &json_value);
    void set_number(const std::string &key, double v);
    void set_bool(const std::string &key, bool v) { set(key, v ? "true" : "false"); }
    void save(const std::string &path) const;
    bool write_back = false;
    std::string loaded_from;

  private:
    std::map<std::string, std::string> kv_; // raw JSON scalar text; strings keep their quotes
};

config_store &get_config();
void set_config(const std::string &key, double v);
void set_config(const std::string &key, const std::string &v);

} // namespace gs
