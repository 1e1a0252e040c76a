// manifest.cc -- host logic either side of the hash: which files belong to a sync, the manifest
// they are recorded in, and which objects a sync must move.
//
// Mirrors, above the C-ABI, what the Go side does around rclone today:
//   - rule assembly of machine.Transfer and LimitTransfer
//       /root/reference/task/common/machine/storage.go:37-41 (built-in excludes), :126-138 (bare
//       paths become "- /path"), :267-284 (LimitTransfer, isRcloneFilter)
//   - [rclone v1.57.0] fs/filter glob rules (first match wins, default include), the local backend's
//     listing (regular files, symlinks skipped), fs/march pairing by name and
//     fs/operations equal()/CheckHashes
//   - the md5sum text manifest of the VM script
//       /root/reference/task/common/machine/machine-script.sh.tpl:111
// Digests are never computed here: tpimd5_manifest_scan hands the file list to tpimd5_hash_files and
// the pairwise digest comparison of tpimd5_plan_transfer goes through tpimd5_diff (both GPU).
#include "../include/tpimd5.h"

#include <algorithm>
#include <cerrno>
#include <cstdio>
#include <cstring>
#include <dirent.h>
#include <string>
#include <condition_variable>
#include <fcntl.h>
#include <mutex>
#include <sys/stat.h>
#include <thread>
#include <vector>

extern "C" void tpimd5_set_error_(const char *msg);   // defined in tpimd5.cu

namespace {

int fail(int code, const std::string &msg) {
    tpimd5_set_error_(msg.c_str());
    return code;
}

// ------------------------------------------------------------------------------------------------
// rclone filter globs
// ------------------------------------------------------------------------------------------------
struct Tok {
    enum Kind { LIT, ANY1, STAR, DSTAR, CLASS } kind;
    char c = 0;
    bool neg = false;
    std::string set;      // CLASS: literal members; ranges expanded lazily via pairs in `ranges`
    std::vector<std::pair<unsigned char, unsigned char>> ranges;
};

struct Pattern {
    bool anchored = false;
    std::vector<Tok> toks;
};

// {a,b} alternatives are expanded into separate globs up front (rclone compiles them into one
// regexp; the language matched is the same).
bool expand_braces(const std::string &g, std::vector<std::string> &out, std::string &err) {
    size_t open = std::string::npos;
    int depth = 0;
    for (size_t i = 0; i < g.size(); ++i) {
        if (g[i] == '\\' && i + 1 < g.size()) { ++i; continue; }
        if (g[i] == '[') {                                   // skip a class
            size_t j = g.find(']', i + 1);
            if (j == std::string::npos) { err = "mismatched '[' in glob"; return false; }
            i = j;
            continue;
        }
        if (g[i] == '{') { if (depth++ == 0) open = i; }
        else if (g[i] == '}') {
            if (depth == 0) { err = "mismatched '}' in glob"; return false; }
            if (--depth == 0) {
                // split g[open+1 .. i) on top-level commas
                std::vector<std::string> alts;
                std::string cur;
                int d = 0;
                for (size_t k = open + 1; k < i; ++k) {
                    if (g[k] == '\\' && k + 1 < i) { cur += g[k]; cur += g[++k]; continue; }
                    if (g[k] == '{') ++d;
                    if (g[k] == '}') --d;
                    if (g[k] == ',' && d == 0) { alts.push_back(cur); cur.clear(); continue; }
                    cur += g[k];
                }
                alts.push_back(cur);
                for (const auto &a : alts) {
                    if (!expand_braces(g.substr(0, open) + a + g.substr(i + 1), out, err)) return false;
                    if (out.size() > 4096) { err = "too many {} alternatives"; return false; }
                }
                return true;
            }
        }
    }
    if (depth) { err = "mismatched '{' in glob"; return false; }
    out.push_back(g);
    return true;
}

bool compile_glob(std::string g, Pattern &p, std::string &err) {
    p.anchored = !g.empty() && g[0] == '/';
    if (p.anchored) g.erase(0, 1);
    for (size_t i = 0; i < g.size(); ++i) {
        Tok t;
        const char c = g[i];
        if (c == '\\' && i + 1 < g.size()) { t.kind = Tok::LIT; t.c = g[++i]; }
        else if (c == '*') {
            size_t j = i;
            while (j < g.size() && g[j] == '*') ++j;
            t.kind = (j - i >= 2) ? Tok::DSTAR : Tok::STAR;
            i = j - 1;
        } else if (c == '?') { t.kind = Tok::ANY1; }
        else if (c == '[') {
            size_t j = g.find(']', i + 1);
            if (j == std::string::npos) { err = "mismatched '[' in glob"; return false; }
            std::string body = g.substr(i + 1, j - i - 1);
            t.kind = Tok::CLASS;
            size_t k = 0;
            if (k < body.size() && (body[k] == '!' || body[k] == '^')) { t.neg = true; ++k; }
            for (; k < body.size(); ++k) {
                if (k + 2 < body.size() && body[k + 1] == '-') {
                    t.ranges.push_back({(unsigned char)body[k], (unsigned char)body[k + 2]});
                    k += 2;
                } else {
                    t.set += body[k];
                }
            }
            i = j;
        } else { t.kind = Tok::LIT; t.c = c; }
        p.toks.push_back(t);
    }
    return true;
}

bool class_has(const Tok &t, unsigned char ch) {
    bool in = t.set.find((char)ch) != std::string::npos;
    for (const auto &r : t.ranges) in = in || (ch >= r.first && ch <= r.second);
    return in != t.neg;
}

// does toks[ti..] match s[si..] to the end?
bool match_here(const std::vector<Tok> &toks, size_t ti, const char *s, size_t si, size_t n) {
    while (ti < toks.size()) {
        const Tok &t = toks[ti];
        switch (t.kind) {
            case Tok::LIT:
                if (si >= n || s[si] != t.c) return false;
                ++si; ++ti; break;
            case Tok::ANY1:
                if (si >= n || s[si] == '/') return false;
                ++si; ++ti; break;
            case Tok::CLASS:
                if (si >= n || !class_has(t, (unsigned char)s[si])) return false;
                ++si; ++ti; break;
            case Tok::STAR:
                for (size_t k = si;; ++k) {
                    if (match_here(toks, ti + 1, s, k, n)) return true;
                    if (k >= n || s[k] == '/') return false;
                }
            case Tok::DSTAR:
                for (size_t k = si; k <= n; ++k)
                    if (match_here(toks, ti + 1, s, k, n)) return true;
                return false;
        }
    }
    return si == n;
}

bool pattern_matches(const Pattern &p, const char *s, size_t n) {
    if (match_here(p.toks, 0, s, 0, n)) return true;
    if (p.anchored) return false;
    for (size_t i = 0; i < n; ++i)                              // "(^|/)": also after every slash
        if (s[i] == '/' && match_here(p.toks, 0, s, i + 1, n)) return true;
    return false;
}

struct Rule {
    bool include = false;
    std::string text;               // normalised "+ glob" / "- glob"
    std::vector<Pattern> pats;      // one per {} expansion
};

// Go filepath.Join("/", p) == Clean("/" + p)
std::string join_root(const std::string &p) {
    std::vector<std::string> parts;
    std::string cur;
    auto flush = [&] {
        if (cur.empty() || cur == ".") { cur.clear(); return; }
        if (cur == "..") { if (!parts.empty()) parts.pop_back(); cur.clear(); return; }
        parts.push_back(cur);
        cur.clear();
    };
    for (char c : p) { if (c == '/') flush(); else cur += c; }
    flush();
    std::string out = "/";
    for (size_t i = 0; i < parts.size(); ++i) { if (i) out += '/'; out += parts[i]; }
    return out;
}

bool is_rclone_filter(const std::string &r) { return r.rfind("+ ", 0) == 0 || r.rfind("- ", 0) == 0; }

}  // namespace

struct tpimd5_filter {
    std::vector<std::string> texts;   // every rule as added, normalised ("+ glob" / "- glob")
    std::vector<Rule> rules;          // file rules: first match decides whether a file is in
    std::vector<Rule> dir_rules;      // directory rules, matched against "path/to/dir/": whether to descend
};

static bool compile_alts(const std::string &glob, std::vector<Pattern> &out, std::string &err) {
    std::vector<std::string> alts;
    if (!expand_braces(glob, alts, err)) return false;
    for (const auto &a : alts) {
        Pattern p;
        if (!compile_glob(a, p, err)) return false;
        out.push_back(std::move(p));
    }
    return true;
}

// [rclone] fs/filter globToDirGlobs: the directories a file glob could have matches in, as globs to be
// matched against a directory path with a trailing slash.  "/a/b/**" -> "/a/b/**/", "/a/b/", "/a/", "/";
// a glob without any slash could match anywhere -> "/**".
static std::vector<std::string> glob_to_dir_globs(std::string glob) {
    std::vector<std::string> out;
    for (;;) {
        const size_t si = glob.rfind('/'), sj = glob.rfind("**");
        long i = si == std::string::npos ? -1 : (long)si;
        const long j = sj == std::string::npos ? -1 : (long)sj;
        std::string what;
        if (j > i) { i = j; what = "**"; }
        if (i < 0) {
            if (out.empty()) out.push_back("/**");
            break;
        }
        glob = glob.substr(0, (size_t)i);
        const std::string g = glob + what + "/";
        if (out.empty() || out.back() != g) out.push_back(g);
    }
    return out;
}

// [rclone] fs/filter Filter.Add: a glob ending in "/" is a directory rule (an excluded "dir/" means
// "dir/**"); a glob containing "**" is both a file and a directory rule; an include file rule also
// includes the directories that lead to it (an exclude rule cannot rule a directory out, except "*").
static bool add_rule(tpimd5_filter &f, std::string rule, std::string &err) {
    if (!is_rclone_filter(rule)) rule = "- " + join_root(rule);            // storage.go:131-134
    const bool include = rule[0] == '+';
    std::string glob = rule.substr(2);
    while (!glob.empty() && glob[0] == ' ') glob.erase(0, 1);
    while (!glob.empty() && glob.back() == ' ') glob.pop_back();
    if (glob.empty()) { err = "empty filter rule"; return false; }
    bool is_dir = glob.back() == '/', is_file = !is_dir;
    if (is_dir && !include) glob += "**";
    if (glob.find("**") != std::string::npos) is_dir = is_file = true;
    Rule r;
    r.include = include;
    r.text = rule;
    if (!compile_alts(glob, r.pats, err)) return false;
    f.texts.push_back(rule);
    if (is_file) {
        f.rules.push_back(r);
        if (include || glob == "*") {
            for (const std::string &dg : glob_to_dir_globs(glob)) {
                Rule d;
                d.include = include;
                d.text = rule;
                if (!compile_alts(dg, d.pats, err)) return false;
                f.dir_rules.push_back(std::move(d));
            }
        }
    }
    if (is_dir) f.dir_rules.push_back(r);
    return true;
}

extern "C" int tpimd5_filter_new(const char *const *exclude, size_t n_exclude, const char *limit_subdir, tpimd5_filter **out) {
    if (!out || (n_exclude && !exclude)) return fail(TPIMD5_EINVAL, "filter_new: null pointer");
    *out = nullptr;
    tpimd5_filter *f = new tpimd5_filter;
    std::string err;
    static const char *const builtin[] = {"- /main.tf", "- /terraform.tfstate*", "- /.terraform**"};   // storage.go:37-41
    bool ok = true;
    for (const char *b : builtin) ok = ok && add_rule(*f, b, err);
    for (size_t i = 0; ok && i < n_exclude; ++i) {
        if (!exclude[i]) { ok = false; err = "null rule"; break; }
        ok = add_rule(*f, exclude[i], err);
    }
    if (ok && limit_subdir) {                                                 // LimitTransfer, storage.go:267-280
        const std::string dir = join_root(limit_subdir);
        if (dir != "/") {
            ok = add_rule(*f, "+ " + dir, err) && add_rule(*f, "+ " + (dir + "/**"), err) && add_rule(*f, "- /**", err);
        }
    }
    if (!ok) {
        delete f;
        return fail(TPIMD5_EINVAL, "filter_new: " + err);
    }
    *out = f;
    return 0;
}

extern "C" void tpimd5_filter_free(tpimd5_filter *f) { delete f; }
extern "C" size_t tpimd5_filter_rule_count(const tpimd5_filter *f) { return f ? f->texts.size() : 0; }
extern "C" const char *tpimd5_filter_rule(const tpimd5_filter *f, size_t i) {
    return (f && i < f->texts.size()) ? f->texts[i].c_str() : nullptr;
}

extern "C" int tpimd5_filter_include(const tpimd5_filter *f, const char *rel_path) {
    if (!f || !rel_path) return fail(TPIMD5_EINVAL, "filter_include: null pointer");
    while (*rel_path == '/') ++rel_path;
    const size_t n = strlen(rel_path);
    for (const Rule &r : f->rules)
        for (const Pattern &p : r.pats)
            if (pattern_matches(p, rel_path, n)) return r.include ? 1 : 0;
    return 1;
}

// Would rclone descend into (and, with copyEmptySrcDirs, create) this directory?  rel_dir without or with
// a trailing slash; the root is always in.
extern "C" int tpimd5_filter_include_dir(const tpimd5_filter *f, const char *rel_dir) {
    if (!f || !rel_dir) return fail(TPIMD5_EINVAL, "filter_include_dir: null pointer");
    while (*rel_dir == '/') ++rel_dir;
    std::string d(rel_dir);
    while (!d.empty() && d.back() == '/') d.pop_back();
    if (d.empty()) return 1;
    d += '/';
    for (const Rule &r : f->dir_rules)
        for (const Pattern &p : r.pats)
            if (pattern_matches(p, d.c_str(), d.size())) return r.include ? 1 : 0;
    return 1;
}

// ------------------------------------------------------------------------------------------------
// manifest
// ------------------------------------------------------------------------------------------------
struct Entry {
    std::string path;        // relative, '/'-separated
    uint64_t size = 0;
    int64_t mtime_ns = 0;
    uint8_t md5[16] = {};
    uint8_t has_md5 = 0;
};

struct tpimd5_manifest {
    std::vector<Entry> e;    // sorted by path (bytewise)
};

static void sort_entries(tpimd5_manifest &m) {
    std::sort(m.e.begin(), m.e.end(), [](const Entry &a, const Entry &b) { return a.path < b.path; });
}

// Lists a local tree the way [rclone] backend/local + fs/walk do under a filter: regular files only
// (symlinks and specials skipped), a directory is entered only if the directory rules allow it, a file is
// kept only if the file rules do.  Directories are listed by a few threads at once: on a million small
// files the lstat calls cost as much as the hashing.
struct TreeLister {
    std::string root;
    const tpimd5_filter *f = nullptr;
    std::mutex mu;
    std::condition_variable cv;
    std::vector<std::string> queue;      // directories (relative) waiting to be listed
    size_t pending = 0;                  // queued + being listed
    int err = 0;
    std::string err_text;
    std::vector<Entry> files;
    std::vector<std::string> dirs;

    void list_one(const std::string &rel, std::vector<Entry> &my_files, std::vector<std::string> &my_dirs, std::vector<std::string> &sub) {
        const std::string dir = rel.empty() ? root : root + "/" + rel;
        DIR *d = opendir(dir.c_str());
        if (!d) { set_err(dir + ": " + strerror(errno)); return; }
        for (;;) {
            struct dirent *de = readdir(d);
            if (!de) break;
            const char *nm = de->d_name;
            if (!strcmp(nm, ".") || !strcmp(nm, "..")) continue;
            struct stat sb;
            if (fstatat(dirfd(d), nm, &sb, AT_SYMLINK_NOFOLLOW) != 0) { set_err(dir + "/" + nm + ": " + strerror(errno)); break; }
            std::string r = rel.empty() ? std::string(nm) : rel + "/" + nm;
            if (S_ISDIR(sb.st_mode)) {
                if (f && tpimd5_filter_include_dir(f, r.c_str()) != 1) continue;
                my_dirs.push_back(r);
                sub.push_back(std::move(r));
            } else if (S_ISREG(sb.st_mode)) {
                if (f && tpimd5_filter_include(f, r.c_str()) != 1) continue;
                Entry e;
                e.path = std::move(r);
                e.size = (uint64_t)sb.st_size;
                e.mtime_ns = (int64_t)sb.st_mtim.tv_sec * 1000000000ll + sb.st_mtim.tv_nsec;
                my_files.push_back(std::move(e));
            }
        }
        closedir(d);
    }
    void set_err(const std::string &t) {
        std::lock_guard<std::mutex> g(mu);
        if (!err) { err = TPIMD5_EIO; err_text = t; }
    }
    void worker() {
        std::vector<Entry> my_files;
        std::vector<std::string> my_dirs, sub;
        for (;;) {
            std::string rel;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return !queue.empty() || pending == 0; });
                if (queue.empty()) break;
                rel = std::move(queue.back());
                queue.pop_back();
            }
            sub.clear();
            list_one(rel, my_files, my_dirs, sub);
            {
                std::lock_guard<std::mutex> g(mu);
                for (auto &x : sub) queue.push_back(std::move(x));
                pending += sub.size();
                --pending;
            }
            cv.notify_all();
        }
        std::lock_guard<std::mutex> g(mu);
        for (auto &e : my_files) files.push_back(std::move(e));
        for (auto &x : my_dirs) dirs.push_back(std::move(x));
    }
    int run(bool missing_ok) {
        struct stat sb;
        if (stat(root.c_str(), &sb) != 0) {
            if (missing_ok && errno == ENOENT) return 0;
            return fail(TPIMD5_EIO, root + ": " + strerror(errno));
        }
        queue.push_back("");
        pending = 1;
        unsigned hc = std::thread::hardware_concurrency();
        const int nth = (int)std::min<unsigned>(hc ? hc : 4, 16);
        std::vector<std::thread> th;
        for (int t = 1; t < nth; ++t) th.emplace_back([this] { worker(); });
        worker();
        for (auto &t : th) t.join();
        if (err) return fail(err, err_text);
        std::sort(dirs.begin(), dirs.end());
        return 0;
    }
};

extern "C" int tpimd5_manifest_scan(tpimd5_ctx *ctx, const char *root, const tpimd5_filter *filter, tpimd5_manifest **out) {
    if (!ctx || !root || !out) return fail(TPIMD5_EINVAL, "manifest_scan: null pointer");
    *out = nullptr;
    tpimd5_manifest *m = new tpimd5_manifest;
    std::string r(root);
    while (r.size() > 1 && r.back() == '/') r.pop_back();
    TreeLister lister;
    lister.root = r;
    lister.f = filter;
    int rc = lister.run(false);
    if (rc) { delete m; return rc; }
    m->e = std::move(lister.files);
    sort_entries(*m);
    const size_t n = m->e.size();
    if (n) {
        std::vector<std::string> full(n);
        std::vector<const char *> paths(n);
        std::vector<uint64_t> sizes(n);
        std::vector<uint8_t> dig(n * 16);
        for (size_t i = 0; i < n; ++i) { full[i] = r + "/" + m->e[i].path; paths[i] = full[i].c_str(); sizes[i] = m->e[i].size; }
        size_t failed = (size_t)-1;
        rc = tpimd5_hash_files(ctx, paths.data(), sizes.data(), n, dig.data(), &failed);
        if (rc) { delete m; return rc; }
        for (size_t i = 0; i < n; ++i) { memcpy(m->e[i].md5, &dig[16 * i], 16); m->e[i].has_md5 = 1; }
    }
    *out = m;
    return 0;
}

extern "C" int tpimd5_manifest_from_arrays(const char *const *paths, const uint64_t *sizes, const int64_t *mtimes_ns,
                                           const uint8_t *md5s /* 16*n or NULL */, const uint8_t *has_md5 /* n or NULL */, size_t n,
                                           tpimd5_manifest **out) {
    if (!out || (n && (!paths || !sizes))) return fail(TPIMD5_EINVAL, "manifest_from_arrays: null pointer");
    tpimd5_manifest *m = new tpimd5_manifest;
    m->e.resize(n);
    for (size_t i = 0; i < n; ++i) {
        if (!paths[i]) { delete m; return fail(TPIMD5_EINVAL, "manifest_from_arrays: null path"); }
        Entry &e = m->e[i];
        e.path = paths[i];
        while (!e.path.empty() && e.path[0] == '/') e.path.erase(0, 1);
        e.size = sizes[i];
        e.mtime_ns = mtimes_ns ? mtimes_ns[i] : 0;
        if (md5s && (!has_md5 || has_md5[i])) { memcpy(e.md5, md5s + 16 * i, 16); e.has_md5 = 1; }
    }
    sort_entries(*m);
    for (size_t i = 1; i < n; ++i)
        if (m->e[i].path == m->e[i - 1].path) {
            std::string dup = m->e[i].path;
            delete m;
            return fail(TPIMD5_EINVAL, "manifest_from_arrays: duplicate path " + dup);
        }
    *out = m;
    return 0;
}

extern "C" void tpimd5_manifest_free(tpimd5_manifest *m) { delete m; }
extern "C" size_t tpimd5_manifest_count(const tpimd5_manifest *m) { return m ? m->e.size() : 0; }

extern "C" int tpimd5_manifest_entry(const tpimd5_manifest *m, size_t i, const char **path, uint64_t *size, int64_t *mtime_ns,
                                     uint8_t md5_out[16], int *has_md5) {
    if (!m || i >= m->e.size()) return fail(TPIMD5_EINVAL, "manifest_entry: index out of range");
    const Entry &e = m->e[i];
    if (path) *path = e.path.c_str();
    if (size) *size = e.size;
    if (mtime_ns) *mtime_ns = e.mtime_ns;
    if (md5_out) memcpy(md5_out, e.md5, 16);
    if (has_md5) *has_md5 = e.has_md5;
    return 0;
}

// Persisted form: "TPIM1\n", u64 count, then per entry u32 path length, path, u64 size, i64 mtime_ns,
// u8 has_md5, 16 bytes md5.  Little-endian, as is (the library targets x86-64/aarch64 hosts only).
extern "C" int tpimd5_manifest_save(const tpimd5_manifest *m, const char *file) {
    if (!m || !file) return fail(TPIMD5_EINVAL, "manifest_save: null pointer");
    FILE *fp = fopen(file, "wb");
    if (!fp) return fail(TPIMD5_EIO, std::string(file) + ": " + strerror(errno));
    bool ok = fwrite("TPIM1\n", 1, 6, fp) == 6;
    const uint64_t n = m->e.size();
    ok = ok && fwrite(&n, 8, 1, fp) == 1;
    for (const Entry &e : m->e) {
        const uint32_t pl = (uint32_t)e.path.size();
        ok = ok && fwrite(&pl, 4, 1, fp) == 1 && fwrite(e.path.data(), 1, pl, fp) == pl && fwrite(&e.size, 8, 1, fp) == 1 &&
             fwrite(&e.mtime_ns, 8, 1, fp) == 1 && fwrite(&e.has_md5, 1, 1, fp) == 1 && fwrite(e.md5, 1, 16, fp) == 16;
    }
    ok = (fclose(fp) == 0) && ok;
    return ok ? 0 : fail(TPIMD5_EIO, std::string(file) + ": write failed");
}

extern "C" int tpimd5_manifest_load(const char *file, tpimd5_manifest **out) {
    if (!file || !out) return fail(TPIMD5_EINVAL, "manifest_load: null pointer");
    *out = nullptr;
    FILE *fp = fopen(file, "rb");
    if (!fp) return fail(TPIMD5_EIO, std::string(file) + ": " + strerror(errno));
    char magic[6];
    uint64_t n = 0;
    bool ok = fread(magic, 1, 6, fp) == 6 && !memcmp(magic, "TPIM1\n", 6) && fread(&n, 8, 1, fp) == 1 && n <= ((uint64_t)1 << 32);
    tpimd5_manifest *m = new tpimd5_manifest;
    if (ok) m->e.resize((size_t)n);
    for (size_t i = 0; ok && i < n; ++i) {
        Entry &e = m->e[i];
        uint32_t pl = 0;
        ok = fread(&pl, 4, 1, fp) == 1 && pl <= (1u << 16);
        if (ok) { e.path.resize(pl); ok = fread(&e.path[0], 1, pl, fp) == pl; }
        ok = ok && fread(&e.size, 8, 1, fp) == 1 && fread(&e.mtime_ns, 8, 1, fp) == 1 && fread(&e.has_md5, 1, 1, fp) == 1 &&
             fread(e.md5, 1, 16, fp) == 16;
    }
    fclose(fp);
    if (!ok) { delete m; return fail(TPIMD5_EIO, std::string(file) + ": not a TPIM1 manifest or truncated"); }
    sort_entries(*m);
    *out = m;
    return 0;
}

// coreutils md5sum text lines: "<32 hex>  <path>\n"; names with backslash or newline use the
// escaped form md5sum itself prints ("\" prefix, "\\" and "\n").
extern "C" int tpimd5_manifest_write_md5sum(const tpimd5_manifest *m, const char *file) {
    if (!m || !file) return fail(TPIMD5_EINVAL, "manifest_write_md5sum: null pointer");
    FILE *fp = fopen(file, "wb");
    if (!fp) return fail(TPIMD5_EIO, std::string(file) + ": " + strerror(errno));
    static const char *hx = "0123456789abcdef";
    bool ok = true;
    for (const Entry &e : m->e) {
        if (!e.has_md5) continue;
        std::string line;
        const bool esc = e.path.find_first_of("\\\n\r") != std::string::npos;
        if (esc) line += '\\';
        for (int i = 0; i < 16; ++i) { line += hx[e.md5[i] >> 4]; line += hx[e.md5[i] & 15]; }
        line += "  ";
        if (!esc) line += e.path;
        else
            for (char c : e.path) {
                if (c == '\\') line += "\\\\";
                else if (c == '\n') line += "\\n";
                else if (c == '\r') line += "\\r";
                else line += c;
            }
        line += '\n';
        ok = ok && fwrite(line.data(), 1, line.size(), fp) == line.size();
    }
    ok = (fclose(fp) == 0) && ok;
    return ok ? 0 : fail(TPIMD5_EIO, std::string(file) + ": write failed");
}

// ------------------------------------------------------------------------------------------------
// which objects must a sync move?  march pairing by name + rclone equal()
// ------------------------------------------------------------------------------------------------
extern "C" int tpimd5_plan_transfer(tpimd5_ctx *ctx, const tpimd5_manifest *src, const tpimd5_manifest *dst, int mode,
                                    int64_t window_ns, uint8_t *actions_out, size_t *n_transfer_out, size_t *n_dst_only_out) {
    if (!src || !dst || (!actions_out && !src->e.empty())) return fail(TPIMD5_EINVAL, "plan_transfer: null pointer");
    if (mode < 0 || mode > 2) return fail(TPIMD5_EINVAL, "plan_transfer: mode 0 (modtime), 1 (checksum) or 2 (size only)");
    const size_t ns = src->e.size(), nd = dst->e.size();
    // pair by name (both sorted)
    std::vector<size_t> partner(ns, (size_t)-1);
    size_t i = 0, j = 0, dst_only = 0;
    while (i < ns && j < nd) {
        const int c = src->e[i].path.compare(dst->e[j].path);
        if (c == 0) { partner[i] = j; ++i; ++j; }
        else if (c < 0) ++i;
        else { ++j; ++dst_only; }
    }
    dst_only += nd - j;
    // pairs whose outcome hangs on the digests: same size, both digests known, and the mode asks
    std::vector<size_t> need;
    for (i = 0; i < ns; ++i) {
        if (partner[i] == (size_t)-1) { actions_out[i] = TPIMD5_ACT_TRANSFER_NEW; continue; }
        const Entry &s = src->e[i], &d = dst->e[partner[i]];
        if (s.size != d.size) { actions_out[i] = TPIMD5_ACT_TRANSFER; continue; }
        if (mode == 2) { actions_out[i] = TPIMD5_ACT_SKIP; continue; }
        const bool both = s.has_md5 && d.has_md5;
        if (mode == 1) {
            if (!both) { actions_out[i] = TPIMD5_ACT_SKIP; continue; }      // unavailable hash: equal by size
            need.push_back(i);
            continue;
        }
        int64_t dt = s.mtime_ns - d.mtime_ns;
        if (dt < 0) dt = -dt;
        if (dt < window_ns || dt == 0) { actions_out[i] = TPIMD5_ACT_SKIP; continue; }
        if (!both) { actions_out[i] = TPIMD5_ACT_TRANSFER; continue; }      // mtime differs, nothing vouches
        need.push_back(i);
    }
    if (!need.empty()) {
        if (!ctx) return fail(TPIMD5_ENODEV, "plan_transfer: digests must be compared and there is no GPU context (no CPU fallback)");
        const size_t k = need.size();
        std::vector<uint8_t> a(16 * k), b(16 * k);
        for (size_t t = 0; t < k; ++t) {
            memcpy(&a[16 * t], src->e[need[t]].md5, 16);
            memcpy(&b[16 * t], dst->e[partner[need[t]]].md5, 16);
        }
        std::vector<uint32_t> changed(k);
        size_t nc = 0;
        int rc = tpimd5_diff(ctx, a.data(), b.data(), k, changed.data(), &nc);
        if (rc) return rc;
        for (size_t t = 0; t < k; ++t) actions_out[need[t]] = (mode == 1) ? TPIMD5_ACT_SKIP : TPIMD5_ACT_FIX_MTIME;
        for (size_t t = 0; t < nc; ++t) actions_out[need[changed[t]]] = TPIMD5_ACT_TRANSFER;
    }
    size_t nt = 0;
    for (i = 0; i < ns; ++i) nt += (actions_out[i] == TPIMD5_ACT_TRANSFER || actions_out[i] == TPIMD5_ACT_TRANSFER_NEW);
    if (n_transfer_out) *n_transfer_out = nt;
    if (n_dst_only_out) *n_dst_only_out = dst_only;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// machine.Transfer for two local directories
// ------------------------------------------------------------------------------------------------
// The reference's Transfer (/root/reference/task/common/machine/storage.go:123-159) hands any pair of
// remotes to rclone's sync.CopyDir(dst, src, copyEmptySrcDirs=true).  Cloud remotes are rclone's
// networking and stay rclone's; the local->local case -- the one storage_test.go:55-101 runs and
// the tail of the Kubernetes Pull (task/k8s/task.go:294-299) -- is reproduced here end to end with
// every digest computed on the GPU: list both sides under the same filter, hash only what the
// comparison mode needs (as rclone does lazily), decide per object with equal(), create the
// directories, copy, carry the modification time over, then verify each copy by digest.
#include <fcntl.h>
#include <atomic>
#include <sys/sendfile.h>
#include <unistd.h>

namespace {

int mkdir_p(const std::string &path) {
    std::string cur;
    for (size_t i = 0; i <= path.size(); ++i) {
        if (i == path.size() || path[i] == '/') {
            if (!cur.empty() && mkdir(cur.c_str(), 0777) != 0 && errno != EEXIST) return fail(TPIMD5_EIO, cur + ": " + strerror(errno));
        }
        if (i < path.size()) cur += path[i];
    }
    return 0;
}

int copy_file(const std::string &from, const std::string &to, int64_t mtime_ns, uint64_t *bytes) {
    int in = open(from.c_str(), O_RDONLY | O_CLOEXEC);
    if (in < 0) return fail(TPIMD5_EIO, from + ": " + strerror(errno));
    int out = open(to.c_str(), O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0666);
    if (out < 0) { int e = errno; close(in); return fail(TPIMD5_EIO, to + ": " + strerror(e)); }
    uint64_t total = 0;
    for (;;) {
        ssize_t k = sendfile(out, in, nullptr, 1u << 30);
        if (k < 0) {
            if (errno == EINTR) continue;
            int e = errno; close(in); close(out);
            return fail(TPIMD5_EIO, to + ": " + strerror(e));
        }
        if (k == 0) break;
        total += (uint64_t)k;
    }
    struct timespec ts[2];
    ts[0].tv_sec = 0; ts[0].tv_nsec = UTIME_OMIT;
    ts[1].tv_sec = mtime_ns / 1000000000ll; ts[1].tv_nsec = mtime_ns % 1000000000ll;
    futimens(out, ts);
    close(in);
    if (close(out) != 0) return fail(TPIMD5_EIO, to + ": " + strerror(errno));
    if (bytes) *bytes = total;
    return 0;
}

std::string strip_slashes(const char *p) {
    std::string r(p);
    while (r.size() > 1 && r.back() == '/') r.pop_back();
    return r;
}

}  // namespace

extern "C" int tpimd5_transfer_local(tpimd5_ctx *ctx, const char *source, const char *destination, const char *const *exclude,
                                     size_t n_exclude, const char *limit_subdir, int mode, tpimd5_transfer_stats *stats_out) {
    tpimd5_transfer_stats st;
    memset(&st, 0, sizeof st);
    if (stats_out) *stats_out = st;
    if (!ctx || !source || !destination) return fail(TPIMD5_EINVAL, "transfer_local: null pointer");
    if (mode < 0 || mode > 2) return fail(TPIMD5_EINVAL, "transfer_local: mode 0 (modtime), 1 (checksum) or 2 (size only)");
    if (source[0] == ':' || destination[0] == ':')
        return fail(TPIMD5_EINVAL, "transfer_local: rclone connection strings (cloud remotes) are moved by rclone itself; "
                                   "this entry point syncs local directories");
    tpimd5_filter *flt = nullptr;
    int rc = tpimd5_filter_new(exclude, n_exclude, limit_subdir, &flt);
    if (rc) return rc;
    const std::string src = strip_slashes(source), dst = strip_slashes(destination);
    TreeLister ls, ld;
    ls.root = src; ls.f = flt;
    ld.root = dst; ld.f = flt;
    rc = ls.run(false);
    if (!rc) rc = ld.run(true);
    tpimd5_filter_free(flt);
    if (rc) return rc;
    tpimd5_manifest ms, md;
    ms.e = std::move(ls.files);
    md.e = std::move(ld.files);
    sort_entries(ms);
    sort_entries(md);
    st.files_listed = ms.e.size();
    for (const Entry &e : ms.e) st.bytes_listed += e.size;                 // operations.Count (storage.go:150)

    // hash lazily, like rclone: only pairs of equal size whose mode consults the digest
    {
        std::vector<std::string> full;
        std::vector<uint64_t> sizes;
        std::vector<Entry *> who;
        size_t i = 0, j = 0;
        while (i < ms.e.size() && j < md.e.size()) {
            const int c = ms.e[i].path.compare(md.e[j].path);
            if (c < 0) { ++i; continue; }
            if (c > 0) { ++j; continue; }
            Entry &a = ms.e[i], &b = md.e[j];
            int64_t dt = a.mtime_ns - b.mtime_ns;
            if (dt < 0) dt = -dt;
            const bool ask = a.size == b.size && (mode == 1 || (mode == 0 && dt >= 1));
            if (ask) {
                full.push_back(src + "/" + a.path); sizes.push_back(a.size); who.push_back(&a);
                full.push_back(dst + "/" + b.path); sizes.push_back(b.size); who.push_back(&b);
            }
            ++i; ++j;
        }
        if (!full.empty()) {
            std::vector<const char *> ptrs(full.size());
            for (size_t k = 0; k < full.size(); ++k) ptrs[k] = full[k].c_str();
            std::vector<uint8_t> dig(16 * full.size());
            rc = tpimd5_hash_files(ctx, ptrs.data(), sizes.data(), full.size(), dig.data(), nullptr);
            if (rc) return rc;
            for (size_t k = 0; k < full.size(); ++k) { memcpy(who[k]->md5, &dig[16 * k], 16); who[k]->has_md5 = 1; st.bytes_hashed += sizes[k]; }
        }
    }
    std::vector<uint8_t> act(ms.e.size() + 1);
    size_t nt = 0, dst_only = 0;
    rc = tpimd5_plan_transfer(ctx, &ms, &md, mode, 1, act.data(), &nt, &dst_only);
    if (rc) return rc;
    st.dst_only = dst_only;

    // directories first (copyEmptySrcDirs = true), then the objects
    rc = mkdir_p(dst);
    for (size_t k = 0; !rc && k < ls.dirs.size(); ++k) rc = mkdir_p(dst + "/" + ls.dirs[k]);
    if (rc) return rc;
    std::vector<size_t> moved;
    for (size_t k = 0; k < ms.e.size(); ++k) {
        const Entry &e = ms.e[k];
        const std::string to = dst + "/" + e.path;
        if (act[k] == TPIMD5_ACT_SKIP) { ++st.skipped; continue; }
        if (act[k] == TPIMD5_ACT_FIX_MTIME) {
            struct timespec ts[2];
            ts[0].tv_sec = 0; ts[0].tv_nsec = UTIME_OMIT;
            ts[1].tv_sec = e.mtime_ns / 1000000000ll; ts[1].tv_nsec = e.mtime_ns % 1000000000ll;
            if (utimensat(AT_FDCWD, to.c_str(), ts, 0) != 0) return fail(TPIMD5_EIO, to + ": " + strerror(errno));
            ++st.mtime_fixed;
            continue;
        }
        const size_t slash = to.rfind('/');
        if (slash != std::string::npos) { rc = mkdir_p(to.substr(0, slash)); if (rc) return rc; }
        moved.push_back(k);
    }
    // the copies themselves, several at a time (rclone: --transfers, 4 by default; local copies are cheap: 8)
    if (!moved.empty()) {
        std::atomic<size_t> next{0};
        std::atomic<uint64_t> copied{0};
        std::mutex err_mu;
        int first_rc = 0;
        std::string first_err;
        auto worker = [&] {
            for (;;) {
                const size_t i = next.fetch_add(1);
                if (i >= moved.size()) break;
                const Entry &e = ms.e[moved[i]];
                uint64_t nbytes = 0;
                const int r = copy_file(src + "/" + e.path, dst + "/" + e.path, e.mtime_ns, &nbytes);
                if (r) {
                    std::lock_guard<std::mutex> g(err_mu);
                    if (!first_rc) { first_rc = r; first_err = tpimd5_last_error(); }
                    next.store(moved.size());               // stop handing out work
                    break;
                }
                copied.fetch_add(nbytes);
            }
        };
        std::vector<std::thread> th;
        const size_t nth = std::min<size_t>(8, moved.size());
        for (size_t t = 1; t < nth; ++t) th.emplace_back(worker);
        worker();
        for (auto &t : th) t.join();
        if (first_rc) return fail(first_rc, first_err);
        st.transferred += moved.size();
        st.bytes_transferred += copied.load();
    }
    // post-transfer check (rclone: size, then the common hash): source and copy hashed in one GPU pass
    if (!moved.empty()) {
        const size_t m = moved.size();
        std::vector<std::string> full(2 * m);
        std::vector<const char *> ptrs(2 * m);
        std::vector<uint64_t> sizes(2 * m);
        for (size_t k = 0; k < m; ++k) {
            full[k] = src + "/" + ms.e[moved[k]].path;
            full[m + k] = dst + "/" + ms.e[moved[k]].path;
            sizes[k] = sizes[m + k] = ms.e[moved[k]].size;
        }
        for (size_t k = 0; k < 2 * m; ++k) ptrs[k] = full[k].c_str();
        std::vector<uint8_t> dig(32 * m);
        size_t failed = (size_t)-1;
        rc = tpimd5_hash_files(ctx, ptrs.data(), sizes.data(), 2 * m, dig.data(), &failed);
        if (rc) return rc;
        std::vector<uint32_t> bad(m);
        size_t nbad = 0;
        rc = tpimd5_diff(ctx, dig.data(), dig.data() + 16 * m, m, bad.data(), &nbad);
        if (rc) return rc;
        for (size_t k = 0; k < m; ++k) st.bytes_hashed += 2 * sizes[k];
        if (nbad) {
            const std::string victim = full[m + bad[0]];
            for (size_t k = 0; k < nbad; ++k) unlink(full[m + bad[k]].c_str());
            return fail(TPIMD5_EIO, victim + ": corrupted on transfer: MD5 hash differ (" + std::to_string(nbad) + " object(s) removed)");
        }
    }
    if (stats_out) *stats_out = st;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// string forms of a digest that the S3 / Azure / GCS back-ends exchange
// ------------------------------------------------------------------------------------------------
// [rclone v1.57.0] backend/s3: Content-MD5 and the Md5chksum metadata are base64 of the 16 bytes;
// Object.Hash accepts the ETag as an MD5 only if it is exactly 32 lowercase hex digits (multipart
// "-N" and SSE-KMS ETags are not), see SURVEY.md appendix C.  azureblob Content-MD5 and GCS md5Hash
// are base64 too; rclone's hash strings are lowercase hex.
extern "C" void tpimd5_hex(const uint8_t md5[16], char out33[33]) {
    static const char *hx = "0123456789abcdef";
    for (int i = 0; i < 16; ++i) { out33[2 * i] = hx[md5[i] >> 4]; out33[2 * i + 1] = hx[md5[i] & 15]; }
    out33[32] = 0;
}

extern "C" void tpimd5_base64(const uint8_t md5[16], char out25[25]) {
    static const char *b64 = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
    int o = 0;
    for (int i = 0; i < 15; i += 3) {
        const uint32_t v = (md5[i] << 16) | (md5[i + 1] << 8) | md5[i + 2];
        out25[o++] = b64[v >> 18]; out25[o++] = b64[(v >> 12) & 63]; out25[o++] = b64[(v >> 6) & 63]; out25[o++] = b64[v & 63];
    }
    const uint32_t v = md5[15] << 16;
    out25[o++] = b64[v >> 18]; out25[o++] = b64[(v >> 12) & 63]; out25[o++] = '='; out25[o++] = '=';
    out25[o] = 0;
}

// 1 and the digest if `etag` (quotes allowed, as S3 returns it) is a plain MD5; 0 if it is a multipart
// or otherwise opaque ETag -- then rclone falls back to the Md5chksum metadata or reports no hash.
// [rclone v1.57.0, restated from memory, not verifiable on this box] backend/s3 lowercases the ETag and
// strips the quotes (strings.Trim(strings.ToLower(etag), `"`)) before matching ^[0-9a-f]{32}$, so an
// uppercase ETag from an S3-compatible store is still an MD5.
extern "C" int tpimd5_md5_from_etag(const char *etag, uint8_t md5_out[16]) {
    if (!etag) return 0;
    std::string e(etag);
    for (char &c : e)
        if (c >= 'A' && c <= 'Z') c = (char)(c - 'A' + 'a');
    size_t lo = 0, hi = e.size();
    while (lo < hi && e[lo] == '"') ++lo;                 // strings.Trim: every leading / trailing quote
    while (hi > lo && e[hi - 1] == '"') --hi;
    e = e.substr(lo, hi - lo);
    if (e.size() != 32) return 0;
    uint8_t d[16];
    for (int i = 0; i < 32; ++i) {
        const char c = e[i];
        int v;
        if (c >= '0' && c <= '9') v = c - '0';
        else if (c >= 'a' && c <= 'f') v = c - 'a' + 10;
        else return 0;
        if (i & 1) d[i >> 1] |= (uint8_t)v; else d[i >> 1] = (uint8_t)(v << 4);
    }
    if (md5_out) memcpy(md5_out, d, 16);
    return 1;
}

// "<32 hex>-<nparts>": the multipart ETag rclone expects back from S3 after CompleteMultipartUpload
extern "C" int tpimd5_s3_etag_text(const uint8_t etag_md5[16], uint32_t nparts, char *out, size_t out_len) {
    char hex[33];
    tpimd5_hex(etag_md5, hex);
    const int k = snprintf(out, out_len, "%s-%u", hex, nparts);
    return (k > 0 && (size_t)k < out_len) ? 0 : fail(TPIMD5_EINVAL, "s3_etag_text: buffer too small");
}
