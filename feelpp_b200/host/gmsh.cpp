// gmsh.cpp -- see gmsh.hpp.  Section grammar as the reference reads it (feelfilters/importergmsh.hpp:1004-1103 header,
// :1106-1420 version 2, readFromFileVersion4 for 4.x); element-type node counts as getInfoMSH (feelfilters/gmshenums).
#include "gmsh.hpp"

#include <algorithm>
#include <array>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <stdexcept>
#include <unordered_map>

namespace feelpp_b200 {
namespace {

[[noreturn]] void fail(std::string const& s) { throw std::runtime_error("gmsh reader: " + s); }

// nodes per element type (Gmsh MSH element types): 1 line, 2 triangle, 3 quad, 4 tet, 5 hex, 6 prism, 7 pyramid,
// 8 line3, 9 triangle6, 10 quad9, 11 tet10, 15 point
int nodesOfType(int t)
{
    static const int n[] = {0, 2, 3, 4, 4, 8, 6, 5, 3, 6, 9, 10, 27, 18, 14, 1};
    return (t >= 1 && t <= 15) ? n[t] : 0;
}

// The whole file in memory and a cursor over it (C stdio and strtod/strtoll only: the library is loaded into
// processes that carry their own C++ runtime, so no iostream state is shared or assumed).
struct Cursor {
    std::vector<char> buf;   // file contents + terminating 0
    size_t pos = 0;
    bool ok = true;
    explicit operator bool() const { return ok; }
    bool operator!() const { return !ok; }
    void skipws() { while (pos + 1 < buf.size() && (buf[pos] == ' ' || buf[pos] == '\n' || buf[pos] == '\r' || buf[pos] == '\t')) ++pos; }
    bool atEnd() { skipws(); return pos + 1 >= buf.size(); }
    int get() { return pos + 1 < buf.size() ? (unsigned char)buf[pos++] : -1; }
    bool word(std::string& w)
    {
        skipws();
        const size_t b = pos;
        while (pos + 1 < buf.size() && !(buf[pos] == ' ' || buf[pos] == '\n' || buf[pos] == '\r' || buf[pos] == '\t')) ++pos;
        w.assign(&buf[b], pos - b);
        if (w.empty()) ok = false;
        return !w.empty();
    }
    void line(std::string& w)
    {
        const size_t b = pos;
        while (pos + 1 < buf.size() && buf[pos] != '\n') ++pos;
        w.assign(&buf[b], pos - b);
    }
    Cursor& operator>>(int64_t& v)
    {
        skipws();
        char* e;
        v = std::strtoll(&buf[pos], &e, 10);
        if (e == &buf[pos]) ok = false;
        pos = (size_t)(e - buf.data());
        return *this;
    }
    Cursor& operator>>(int& v) { int64_t t = 0; *this >> t; v = (int)t; return *this; }
    Cursor& operator>>(double& v)
    {
        skipws();
        char* e;
        v = std::strtod(&buf[pos], &e);
        if (e == &buf[pos]) ok = false;
        pos = (size_t)(e - buf.data());
        return *this;
    }
    Cursor& operator>>(std::string& w) { word(w); return *this; }
    void read(char* dst, size_t n)
    {
        if (pos + n + 1 > buf.size()) { ok = false; return; }
        std::memcpy(dst, &buf[pos], n);
        pos += n;
    }
};

void expect(Cursor& is, const char* what)
{
    std::string s;
    if (!is.word(s) || s != what) fail(std::string("expected ") + what + ", found '" + s + "'");
}

template <class T> T rd(Cursor& is, bool swap)
{
    T v;
    is.read(reinterpret_cast<char*>(&v), sizeof(T));
    if (!is) fail("truncated binary data");
    if (swap) std::reverse(reinterpret_cast<char*>(&v), reinterpret_cast<char*>(&v) + sizeof(T));
    return v;
}

struct Raw {
    std::vector<std::pair<int64_t, std::array<double, 3>>> nodes;
    struct Elt { int type; int physical; int elementary; int partition; int64_t v[4]; };
    std::vector<Elt> elts;
};

void keep(Raw& R, int type, int physical, int elementary, int partition, const int64_t* idx)
{
    if (type == 4 || type == 2) {
        Raw::Elt e{type, physical, elementary, partition, {0, 0, 0, 0}};
        for (int j = 0; j < nodesOfType(type); ++j) e.v[j] = idx[j];
        R.elts.push_back(e);
    } else if (type == 1 || type == 15) {
        // lines and points of the geometry: carried by the file, not used by the path
    } else
        fail("element type " + std::to_string(type) + " is not an order-1 tetrahedron / triangle (engine scope: Simplex<3,1>)");
}

void readV2(Cursor& is, bool binary, bool swap, Raw& R)
{
    std::string buf;
    is >> buf;
    if (buf != "$Nodes") fail("expected $Nodes, found '" + buf + "' ($ParametricNodes and $NOD are not supported)");
    int64_t n;
    is >> n;
    if (binary) is.get();
    R.nodes.resize((size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        auto& p = R.nodes[(size_t)i];
        if (!binary) {
            if (!(is >> p.first >> p.second[0] >> p.second[1] >> p.second[2])) fail("bad node line");
        } else {
            p.first = rd<int32_t>(is, swap);
            for (int c = 0; c < 3; ++c) p.second[c] = rd<double>(is, swap);
        }
    }
    if (binary) is.get();
    expect(is, "$EndNodes");
    expect(is, "$Elements");
    int64_t ne;
    is >> ne;
    if (binary) is.get();
    if (!binary) {
        for (int64_t i = 0; i < ne; ++i) {
            int64_t num;
            int type, ntags;
            if (!(is >> num >> type >> ntags)) fail("bad element line");
            int physical = 0, elementary = 0, partition = 0;
            for (int j = 0; j < ntags; ++j) {
                int tag;
                is >> tag;
                if (j == 0) physical = tag;
                else if (j == 1) elementary = tag;
                else if (j == 3 && ntags > 3) partition = tag - 1;     // tag 2 = number of partitions, then the owner, then -ghosts
            }
            const int nv = nodesOfType(type);
            if (!nv) fail("unknown number of vertices for element type " + std::to_string(type));
            int64_t idx[27];
            for (int j = 0; j < nv; ++j) is >> idx[j];
            if (!is) fail("bad element line");
            keep(R, type, physical, elementary, partition, idx);
        }
    } else {
        int64_t done = 0;
        while (done < ne) {
            const int type = rd<int32_t>(is, swap), cnt = rd<int32_t>(is, swap), ntags = rd<int32_t>(is, swap);
            const int nv = nodesOfType(type);
            if (!nv) fail("unknown number of vertices for element type " + std::to_string(type));
            for (int i = 0; i < cnt; ++i) {
                rd<int32_t>(is, swap);   // element number
                int physical = 0, elementary = 0, partition = 0;
                for (int j = 0; j < ntags; ++j) {
                    const int tag = rd<int32_t>(is, swap);
                    if (j == 0) physical = tag;
                    else if (j == 1) elementary = tag;
                    else if (j == 3 && ntags > 3) partition = tag - 1;
                }
                int64_t idx[27];
                for (int j = 0; j < nv; ++j) idx[j] = rd<int32_t>(is, swap);
                keep(R, type, physical, elementary, partition, idx);
            }
            done += cnt;
        }
        is.get();
    }
    expect(is, "$EndElements");
}

// MSH 4.1 binary: the same sections, counts as size_t (8 bytes), tags as int, node / element tags as size_t
void readV4Binary(Cursor& is, bool swap, Raw& R)
{
    std::map<std::pair<int, int>, int> phys;
    std::string buf;
    auto sz = [&]() { return (int64_t)rd<uint64_t>(is, swap); };
    while (!is.atEnd() && is.word(buf)) {
        if (buf == "$Entities") {
            is.get();
            int64_t cnt[4];
            for (int d = 0; d < 4; ++d) cnt[d] = sz();
            for (int dim = 0; dim < 4; ++dim)
                for (int64_t k = 0; k < cnt[dim]; ++k) {
                    const int tag = rd<int32_t>(is, swap);
                    for (int c = 0; c < (dim == 0 ? 3 : 6); ++c) rd<double>(is, swap);
                    const int64_t np = sz();
                    for (int64_t j = 0; j < np; ++j) {
                        const int p = rd<int32_t>(is, swap);
                        if (j == 0) phys[{dim, tag}] = p;
                    }
                    if (dim > 0) {
                        const int64_t nb = sz();
                        for (int64_t j = 0; j < nb; ++j) rd<int32_t>(is, swap);
                    }
                }
            expect(is, "$EndEntities");
        } else if (buf == "$Nodes") {
            is.get();
            const int64_t nblk = sz(), n = sz();
            sz();
            sz();
            R.nodes.reserve((size_t)n);
            for (int64_t b = 0; b < nblk; ++b) {
                const int dim = rd<int32_t>(is, swap);
                rd<int32_t>(is, swap);
                const int parametric = rd<int32_t>(is, swap);
                const int64_t nb = sz();
                const size_t first = R.nodes.size();
                R.nodes.resize(first + (size_t)nb);
                for (int64_t i = 0; i < nb; ++i) R.nodes[first + (size_t)i].first = sz();
                for (int64_t i = 0; i < nb; ++i) {
                    auto& x = R.nodes[first + (size_t)i].second;
                    for (int c = 0; c < 3; ++c) x[c] = rd<double>(is, swap);
                    for (int c = 0; parametric && c < dim; ++c) rd<double>(is, swap);
                }
            }
            if ((int64_t)R.nodes.size() != n) fail("bad $Nodes section");
            expect(is, "$EndNodes");
        } else if (buf == "$Elements") {
            is.get();
            const int64_t nblk = sz();
            sz();
            sz();
            sz();
            for (int64_t b = 0; b < nblk; ++b) {
                const int dim = rd<int32_t>(is, swap), tag = rd<int32_t>(is, swap), type = rd<int32_t>(is, swap);
                const int64_t nb = sz();
                const int nv = nodesOfType(type);
                if (!nv) fail("unknown number of vertices for element type " + std::to_string(type));
                auto it = phys.find({dim, tag});
                const int physical = it == phys.end() ? 0 : it->second;
                for (int64_t i = 0; i < nb; ++i) {
                    int64_t idx[27];
                    sz();
                    for (int j = 0; j < nv; ++j) idx[j] = sz();
                    keep(R, type, physical, tag, 0, idx);
                }
            }
            expect(is, "$EndElements");
            return;
        } else if (buf[0] == '$') {   // a section the path does not use: find its end marker in the bytes
            const std::string end = "\n$End" + buf.substr(1);
            const auto at = std::search(is.buf.begin() + (long)is.pos, is.buf.end(), end.begin(), end.end());
            if (at == is.buf.end()) fail("section " + buf + " has no end");
            is.pos = (size_t)(at - is.buf.begin()) + end.size();
        }
    }
    fail("no $Elements section");
}

void readV4(Cursor& is, Raw& R)
{
    // entity (dim, tag) -> first physical tag
    std::map<std::pair<int, int>, int> phys;
    std::string buf;
    while (!is.atEnd() && is.word(buf)) {
        if (buf == "$Entities") {
            int64_t cnt[4];
            is >> cnt[0] >> cnt[1] >> cnt[2] >> cnt[3];
            for (int dim = 0; dim < 4; ++dim)
                for (int64_t k = 0; k < cnt[dim]; ++k) {
                    int tag;
                    double box[6];
                    is >> tag;
                    for (int c = 0; c < (dim == 0 ? 3 : 6); ++c) is >> box[c];
                    int64_t np;
                    is >> np;
                    for (int64_t j = 0; j < np; ++j) {
                        int p;
                        is >> p;
                        if (j == 0) phys[{dim, tag}] = p;
                    }
                    if (dim > 0) {
                        int64_t nb;
                        is >> nb;
                        for (int64_t j = 0; j < nb; ++j) { int b; is >> b; }
                    }
                }
            if (!is) fail("bad $Entities section");
            expect(is, "$EndEntities");
        } else if (buf == "$Nodes") {
            int64_t nblk, n, mn, mx;
            is >> nblk >> n >> mn >> mx;
            R.nodes.reserve((size_t)n);
            for (int64_t b = 0; b < nblk; ++b) {
                int dim, tag, parametric;
                int64_t nb;
                is >> dim >> tag >> parametric >> nb;
                const size_t first = R.nodes.size();
                R.nodes.resize(first + (size_t)nb);
                for (int64_t i = 0; i < nb; ++i) is >> R.nodes[first + (size_t)i].first;
                for (int64_t i = 0; i < nb; ++i) {
                    auto& x = R.nodes[first + (size_t)i].second;
                    is >> x[0] >> x[1] >> x[2];
                    double skip;
                    for (int c = 0; parametric && c < dim; ++c) is >> skip;
                }
            }
            if (!is || (int64_t)R.nodes.size() != n) fail("bad $Nodes section");
            expect(is, "$EndNodes");
        } else if (buf == "$Elements") {
            int64_t nblk, n, mn, mx;
            is >> nblk >> n >> mn >> mx;
            for (int64_t b = 0; b < nblk; ++b) {
                int dim, tag, type;
                int64_t nb;
                is >> dim >> tag >> type >> nb;
                const int nv = nodesOfType(type);
                if (!nv) fail("unknown number of vertices for element type " + std::to_string(type));
                auto it = phys.find({dim, tag});
                const int physical = it == phys.end() ? 0 : it->second;
                for (int64_t i = 0; i < nb; ++i) {
                    int64_t num, idx[27];
                    is >> num;
                    for (int j = 0; j < nv; ++j) is >> idx[j];
                    if (!is) fail("bad element line");
                    keep(R, type, physical, tag, 0, idx);
                }
            }
            expect(is, "$EndElements");
            return;
        } else if (buf[0] == '$') {   // a section the path does not use: skip to its end
            const std::string end = "$End" + buf.substr(1);
            while (!is.atEnd() && is.word(buf) && buf != end) {}
        }
    }
    fail("no $Elements section");
}

} // namespace

GmshMesh readGmsh(std::string const& path, bool elementaryAsPhysical)
{
    Cursor is;
    {
        std::FILE* fp = std::fopen(path.c_str(), "rb");
        if (!fp) fail("cannot open " + path);
        std::fseek(fp, 0, SEEK_END);
        const long len = std::ftell(fp);
        std::fseek(fp, 0, SEEK_SET);
        is.buf.resize((size_t)std::max(len, 0L) + 1, 0);
        const size_t got = len > 0 ? std::fread(is.buf.data(), 1, (size_t)len, fp) : 0;
        std::fclose(fp);
        if ((long)got != len) fail("cannot read " + path);
    }
    GmshMesh M;
    expect(is, "$MeshFormat");
    int format, size;
    is >> M.version >> format >> size;
    if (!is) fail("bad $MeshFormat line");
    if (M.version < 2) fail("only Gmsh version >= 2 is supported");    // importergmsh.hpp:1014-1015
    // MSH 4.0 lays out $Entities / $Nodes / $Elements differently from 4.1 (importergmsh.hpp:1932-2143): not read here
    if (M.version >= 4 && M.version < 4.1) fail("MSH 4.0 files are not supported (re-save as 4.1 or 2.2)");
    M.binary = format != 0;
    bool swap = false;
    if (M.binary) {
        is.get();
        int one = 0;
        is.read(reinterpret_cast<char*>(&one), sizeof(int));
        swap = one != 1;      // importergmsh.hpp:1019-1033
    }
    expect(is, "$EndMeshFormat");
    Raw R;
    if (M.version < 4) {
        const size_t at = is.pos;
        std::string buf;
        is >> buf;
        if (buf == "$PhysicalNames") {
            int64_t n;
            is >> n;
            for (int64_t i = 0; i < n; ++i) {
                int dim = -1, tag;
                std::string name;
                if (M.version >= 2.1) is >> dim >> tag; else is >> tag;      // :1058-1070
                is.line(name);
                const size_t a = name.find('"'), b = name.rfind('"');
                if (a != std::string::npos && b > a) name = name.substr(a + 1, b - a - 1);
                M.physicalNames += std::to_string(dim) + " " + std::to_string(tag) + " " + name + "\n";
            }
            expect(is, "$EndPhysicalNames");
        } else
            is.pos = at;
        readV2(is, M.binary, swap, R);
    } else {
        const size_t at = is.pos;
        std::string buf;
        is >> buf;
        if (buf == "$PhysicalNames") {
            int64_t n;
            is >> n;
            for (int64_t i = 0; i < n; ++i) {
                int dim, tag;
                std::string name;
                is >> dim >> tag;
                is.line(name);
                const size_t a = name.find('"'), b = name.rfind('"');
                if (a != std::string::npos && b > a) name = name.substr(a + 1, b - a - 1);
                M.physicalNames += std::to_string(dim) + " " + std::to_string(tag) + " " + name + "\n";
            }
            expect(is, "$EndPhysicalNames");
        } else
            is.pos = at;
        if (M.binary) readV4Binary(is, swap, R); else readV4(is, R);
    }
    // compact the node numbers in ascending order (order-preserving: id comparisons are unchanged)
    std::sort(R.nodes.begin(), R.nodes.end(), [](auto const& a, auto const& b) { return a.first < b.first; });
    std::unordered_map<int64_t, int32_t> index;
    index.reserve(R.nodes.size() * 2);
    M.xyz.resize(R.nodes.size() * 3);
    M.pointId.resize(R.nodes.size());
    for (size_t i = 0; i < R.nodes.size(); ++i) {
        if (i && R.nodes[i].first == R.nodes[i - 1].first) fail("node " + std::to_string(R.nodes[i].first) + " appears twice");
        index.emplace(R.nodes[i].first, (int32_t)i);
        M.pointId[i] = R.nodes[i].first;
        for (int c = 0; c < 3; ++c) M.xyz[3 * i + c] = R.nodes[i].second[c];
    }
    for (auto const& e : R.elts) {
        const int nv = e.type == 4 ? 4 : 3;
        int32_t v[4];
        for (int j = 0; j < nv; ++j) {
            auto it = index.find(e.v[j]);
            if (it == index.end()) fail("element refers to unknown node " + std::to_string(e.v[j]));
            v[j] = it->second;
        }
        const int marker = elementaryAsPhysical ? e.elementary : e.physical;
        if (e.type == 4) {
            M.conn.insert(M.conn.end(), v, v + 4);
            M.eltMarker.push_back(marker);
            if (e.partition < 0) fail("negative partition tag");
            M.eltPartition.push_back(e.partition);
            M.numPartitions = std::max(M.numPartitions, e.partition + 1);
        } else {
            M.tri.insert(M.tri.end(), v, v + 3);
            M.triMarker.push_back(marker);
        }
    }
    if (M.conn.empty()) fail("the file holds no tetrahedra");
    return M;
}

} // namespace feelpp_b200
