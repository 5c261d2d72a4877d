// Parameter files for the command-line driver.
//
// The reference's executable fills a Teuchos::ParameterList from an XML file
// (/root/reference/src/main.cpp:55-61, Teuchos::updateParametersFromXmlFile)
// and hands the sublists "Lyapunov Solver" and "Eigenvalue Solver" to the
// solver and to the eigenvalue post-processing (:117,143-151).  The solver
// only needs `template<class T> T get(std::string const &, T)`
// (src/LyapunovSolver.hpp:40-70).  ParameterTree provides that, sublists, and
// readers for two file formats:
//
//   * the Teuchos XML subset the reference uses:
//       <ParameterList name="...">
//         <ParameterList name="Lyapunov Solver">
//           <Parameter name="Tolerance" type="double" value="1e-4"/>
//         </ParameterList>
//       </ParameterList>
//   * an INI file: `[Lyapunov Solver]` section headers, `key = value` lines,
//     `#` or `;` comments.
//
// Values are kept as text and converted on get(), so `value="1e-4"` can be
// read as double, `value="true"` as bool, and a double-typed "10" as int.
#ifndef RAILS_B200_PARAMETERFILE_HPP
#define RAILS_B200_PARAMETERFILE_HPP

#include <map>
#include <string>
#include <vector>

namespace RAILS
{

class ParameterTree
{
    std::string name_;
    std::map<std::string, std::string> values_;
    std::map<std::string, ParameterTree> sublists_;

    static bool convert(std::string const &text, bool &out);
    static bool convert(std::string const &text, int &out);
    static bool convert(std::string const &text, double &out);
    static bool convert(std::string const &text, std::string &out);

public:
    ParameterTree() {}
    explicit ParameterTree(std::string const &name) : name_(name) {}

    std::string const &name() const { return name_; }

    // value of `name`, or `def` when it is absent or does not parse as T
    template<class T>
    T get(std::string const &name, T def) const
    {
        std::map<std::string, std::string>::const_iterator it = values_.find(name);
        if (it == values_.end())
            return def;
        T out = def;
        return convert(it->second, out) ? out : def;
    }
    std::string get(std::string const &name, const char *def) const
    { return get(name, std::string(def)); }

    void set(std::string const &name, std::string const &value) { values_[name] = value; }
    void set(std::string const &name, const char *value) { values_[name] = value; }
    void set(std::string const &name, double value);
    void set(std::string const &name, int value);
    void set(std::string const &name, bool value);

    bool isParameter(std::string const &name) const { return values_.count(name) != 0; }
    bool isSublist(std::string const &name) const { return sublists_.count(name) != 0; }
    // created when missing, like Teuchos::ParameterList::sublist
    ParameterTree &sublist(std::string const &name);
    ParameterTree const *find_sublist(std::string const &name) const;

    std::vector<std::string> keys() const;
    std::vector<std::string> sublist_names() const;

    // 0 on success; on failure non-zero and a one-line message in err
    static int from_xml(std::string const &text, ParameterTree &out, std::string &err);
    static int from_ini(std::string const &text, ParameterTree &out, std::string &err);
    // XML when the first non-blank character is '<', INI otherwise
    static int from_file(std::string const &path, ParameterTree &out, std::string &err);

    // Teuchos-style XML text of this list (round-trips through from_xml)
    std::string to_xml(int indent = 0) const;
};

}

#endif
