#include "ParameterFile.hpp"

#include <exception>

#include <cctype>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>

namespace RAILS
{

namespace
{

std::string trim(std::string const &s)
{
    size_t a = 0, b = s.size();
    while (a < b && std::isspace((unsigned char)s[a])) ++a;
    while (b > a && std::isspace((unsigned char)s[b - 1])) --b;
    return s.substr(a, b - a);
}

std::string lower(std::string s)
{
    for (size_t i = 0; i < s.size(); ++i)
        s[i] = (char)std::tolower((unsigned char)s[i]);
    return s;
}

std::string unescape(std::string const &s)
{
    std::string out;
    for (size_t i = 0; i < s.size(); ++i)
    {
        if (s[i] == '&')
        {
            static const char *const names[] = {"&amp;", "&lt;", "&gt;", "&quot;", "&apos;"};
            static const char chars[] = {'&', '<', '>', '"', '\''};
            bool hit = false;
            for (int k = 0; k < 5; ++k)
                if (s.compare(i, std::strlen(names[k]), names[k]) == 0)
                {
                    out += chars[k];
                    i += std::strlen(names[k]) - 1;
                    hit = true;
                    break;
                }
            if (hit)
                continue;
        }
        out += s[i];
    }
    return out;
}

std::string escape(std::string const &s)
{
    std::string out;
    for (size_t i = 0; i < s.size(); ++i)
        switch (s[i])
        {
        case '&': out += "&amp;"; break;
        case '<': out += "&lt;"; break;
        case '>': out += "&gt;"; break;
        case '"': out += "&quot;"; break;
        default: out += s[i];
        }
    return out;
}

// one XML tag: name, attributes, and whether it opens / closes / is empty
struct Tag
{
    std::string name;
    std::map<std::string, std::string> attr;
    bool closing = false;     // </name>
    bool self_closed = false; // <name ... />
};

// parses the tag that starts at text[pos] == '<'; leaves pos behind '>'
bool parse_tag(std::string const &text, size_t &pos, Tag &tag, std::string &err)
{
    size_t end = pos + 1;
    char quote = 0;
    while (end < text.size() && (quote || text[end] != '>'))
    {
        if (quote)
        {
            if (text[end] == quote) quote = 0;
        }
        else if (text[end] == '"' || text[end] == '\'')
            quote = text[end];
        ++end;
    }
    if (end >= text.size())
    {
        err = "unterminated tag";
        return false;
    }
    std::string body = text.substr(pos + 1, end - pos - 1);
    pos = end + 1;
    tag = Tag();
    size_t i = 0;
    if (!body.empty() && body[0] == '/')
    {
        tag.closing = true;
        i = 1;
    }
    if (!body.empty() && body[body.size() - 1] == '/')
    {
        tag.self_closed = true;
        body.erase(body.size() - 1);
    }
    while (i < body.size() && std::isspace((unsigned char)body[i])) ++i;
    size_t j = i;
    while (j < body.size() && !std::isspace((unsigned char)body[j])) ++j;
    tag.name = body.substr(i, j - i);
    i = j;
    for (;;)
    {
        while (i < body.size() && std::isspace((unsigned char)body[i])) ++i;
        if (i >= body.size())
            break;
        size_t eq = body.find('=', i);
        if (eq == std::string::npos)
        {
            err = "attribute without a value in <" + tag.name + ">";
            return false;
        }
        std::string key = trim(body.substr(i, eq - i));
        i = eq + 1;
        while (i < body.size() && std::isspace((unsigned char)body[i])) ++i;
        if (i >= body.size() || (body[i] != '"' && body[i] != '\''))
        {
            err = "attribute '" + key + "' of <" + tag.name + "> is not quoted";
            return false;
        }
        char q = body[i++];
        size_t close = body.find(q, i);
        if (close == std::string::npos)
        {
            err = "unterminated attribute '" + key + "'";
            return false;
        }
        tag.attr[key] = unescape(body.substr(i, close - i));
        i = close + 1;
    }
    return true;
}

}

bool ParameterTree::convert(std::string const &text, bool &out)
{
    std::string t = lower(trim(text));
    if (t == "true" || t == "1" || t == "yes" || t == "on")
    {
        out = true;
        return true;
    }
    if (t == "false" || t == "0" || t == "no" || t == "off")
    {
        out = false;
        return true;
    }
    double d = 0.0;
    if (convert(text, d))
    {
        out = d != 0.0;
        return true;
    }
    return false;
}

bool ParameterTree::convert(std::string const &text, double &out)
{
    std::string t = trim(text);
    if (t.empty())
        return false;
    std::string l = lower(t);
    if (l == "true" || l == "false")
    {
        out = l == "true" ? 1.0 : 0.0;
        return true;
    }
    char *e = nullptr;
    double v = std::strtod(t.c_str(), &e);
    if (e == t.c_str() || *e != '\0')
        return false;
    out = v;
    return true;
}

bool ParameterTree::convert(std::string const &text, int &out)
{
    double d = 0.0;
    if (!convert(text, d))
        return false;
    out = (int)d;
    return true;
}

bool ParameterTree::convert(std::string const &text, std::string &out)
{
    out = text;
    return true;
}

void ParameterTree::set(std::string const &name, double value)
{
    char buf[64];
    std::snprintf(buf, sizeof buf, "%.17g", value);
    values_[name] = buf;
}

void ParameterTree::set(std::string const &name, int value)
{
    values_[name] = std::to_string(value);
}

void ParameterTree::set(std::string const &name, bool value)
{
    values_[name] = value ? "true" : "false";
}

ParameterTree &ParameterTree::sublist(std::string const &name)
{
    std::map<std::string, ParameterTree>::iterator it = sublists_.find(name);
    if (it == sublists_.end())
        it = sublists_.insert(std::make_pair(name, ParameterTree(name))).first;
    return it->second;
}

ParameterTree const *ParameterTree::find_sublist(std::string const &name) const
{
    std::map<std::string, ParameterTree>::const_iterator it = sublists_.find(name);
    return it == sublists_.end() ? nullptr : &it->second;
}

std::vector<std::string> ParameterTree::keys() const
{
    std::vector<std::string> out;
    for (std::map<std::string, std::string>::const_iterator it = values_.begin(); it != values_.end(); ++it)
        out.push_back(it->first);
    return out;
}

std::vector<std::string> ParameterTree::sublist_names() const
{
    std::vector<std::string> out;
    for (std::map<std::string, ParameterTree>::const_iterator it = sublists_.begin(); it != sublists_.end(); ++it)
        out.push_back(it->first);
    return out;
}

int ParameterTree::from_xml(std::string const &text, ParameterTree &out, std::string &err)
{
    std::vector<ParameterTree *> stack;
    bool have_root = false;
    size_t pos = 0;
    while (pos < text.size())
    {
        size_t lt = text.find('<', pos);
        if (lt == std::string::npos)
            break;
        if (text.compare(lt, 4, "<!--") == 0)
        {
            size_t e = text.find("-->", lt + 4);
            if (e == std::string::npos)
            {
                err = "unterminated comment";
                return 1;
            }
            pos = e + 3;
            continue;
        }
        if (text.compare(lt, 2, "<?") == 0)
        {
            size_t e = text.find("?>", lt + 2);
            if (e == std::string::npos)
            {
                err = "unterminated processing instruction";
                return 1;
            }
            pos = e + 2;
            continue;
        }
        if (text.compare(lt, 2, "<!") == 0)
        {
            size_t e = text.find('>', lt);
            if (e == std::string::npos)
            {
                err = "unterminated declaration";
                return 1;
            }
            pos = e + 1;
            continue;
        }
        pos = lt;
        Tag tag;
        if (!parse_tag(text, pos, tag, err))
            return 1;
        if (tag.name == "ParameterList")
        {
            if (tag.closing)
            {
                if (stack.empty())
                {
                    err = "</ParameterList> without an open list";
                    return 1;
                }
                stack.pop_back();
                continue;
            }
            ParameterTree *list = nullptr;
            if (stack.empty())
            {
                if (have_root)
                {
                    err = "more than one top-level ParameterList";
                    return 1;
                }
                have_root = true;
                out = ParameterTree(tag.attr.count("name") ? tag.attr["name"] : std::string());
                list = &out;
            }
            else
            {
                if (!tag.attr.count("name"))
                {
                    err = "nested ParameterList without a name";
                    return 1;
                }
                list = &stack.back()->sublist(tag.attr["name"]);
            }
            if (!tag.self_closed)
                stack.push_back(list);
        }
        else if (tag.name == "Parameter")
        {
            if (tag.closing)
                continue;
            if (stack.empty())
            {
                err = "Parameter outside a ParameterList";
                return 1;
            }
            if (!tag.attr.count("name") || !tag.attr.count("value"))
            {
                err = "Parameter needs the attributes name and value";
                return 1;
            }
            stack.back()->set(tag.attr["name"], tag.attr["value"]);
        }
        else
        {
            err = "unexpected element <" + tag.name + ">";
            return 1;
        }
    }
    if (!stack.empty())
    {
        err = "ParameterList '" + stack.back()->name() + "' is not closed";
        return 1;
    }
    if (!have_root)
    {
        err = "no ParameterList element found";
        return 1;
    }
    return 0;
}

int ParameterTree::from_ini(std::string const &text, ParameterTree &out, std::string &err)
{
    out = ParameterTree();
    ParameterTree *current = &out;
    std::istringstream is(text);
    std::string line;
    int lineno = 0;
    while (std::getline(is, line))
    {
        ++lineno;
        std::string t = trim(line);
        if (t.empty() || t[0] == '#' || t[0] == ';')
            continue;
        if (t[0] == '[')
        {
            size_t close = t.find(']');
            if (close == std::string::npos)
            {
                err = "line " + std::to_string(lineno) + ": section header without ']'";
                return 1;
            }
            // "[A/B]" and "[A.B]" are not split: Teuchos names contain spaces
            // and dots freely, nesting is one level as in the reference
            current = &out.sublist(trim(t.substr(1, close - 1)));
            continue;
        }
        size_t eq = t.find('=');
        if (eq == std::string::npos)
        {
            err = "line " + std::to_string(lineno) + ": expected `key = value`";
            return 1;
        }
        std::string key = trim(t.substr(0, eq));
        std::string value = trim(t.substr(eq + 1));
        if (value.size() >= 2 && (value[0] == '"' || value[0] == '\'') &&
            value[value.size() - 1] == value[0])
            value = value.substr(1, value.size() - 2);
        if (key.empty())
        {
            err = "line " + std::to_string(lineno) + ": empty key";
            return 1;
        }
        current->set(key, value);
    }
    return 0;
}

int ParameterTree::from_file(std::string const &path, ParameterTree &out, std::string &err)
{
    FILE *f = std::fopen(path.c_str(), "rb");
    if (!f)
    {
        err = "cannot open " + path + ": " + std::strerror(errno);
        return 1;
    }
    std::string text;
    char buf[4096];
    size_t got;
    while ((got = std::fread(buf, 1, sizeof buf, f)) > 0)
        text.append(buf, got);
    std::fclose(f);
    size_t i = 0;
    while (i < text.size() && std::isspace((unsigned char)text[i])) ++i;
    int rc = 1;
    try
    {
        rc = (i < text.size() && text[i] == '<') ? from_xml(text, out, err) : from_ini(text, out, err);
    }
    catch (std::exception const &e)          // no exception leaves the reader
    {
        err = e.what();
    }
    if (rc)
        err = path + ": " + err;
    return rc;
}

std::string ParameterTree::to_xml(int indent) const
{
    std::string pad((size_t)indent, ' ');
    std::ostringstream os;
    os << pad << "<ParameterList name=\"" << escape(name_) << "\">\n";
    for (std::map<std::string, std::string>::const_iterator it = values_.begin(); it != values_.end(); ++it)
    {
        bool b = false;
        double d = 0.0;
        std::string l = lower(trim(it->second));
        const char *type = "string";
        if (l == "true" || l == "false")
            type = "bool";
        else if (convert(it->second, d))
            type = (l.find_first_of(".een") == std::string::npos) ? "int" : "double";
        (void)b;
        os << pad << "  <Parameter name=\"" << escape(it->first) << "\" type=\"" << type
           << "\" value=\"" << escape(it->second) << "\"/>\n";
    }
    for (std::map<std::string, ParameterTree>::const_iterator it = sublists_.begin(); it != sublists_.end(); ++it)
        os << it->second.to_xml(indent + 2);
    os << pad << "</ParameterList>\n";
    return os.str();
}

}
