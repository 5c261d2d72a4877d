#include "SparseLU.hpp"

#include <algorithm>
#include <cmath>
#include <functional>
#include <queue>
#include <utility>

namespace RAILS
{

std::vector<int> minimum_degree_order(int n, const int *rowptr, const int *colidx)
{
    // explicit elimination graph: adj[v] = sorted neighbours still to be eliminated
    std::vector<std::vector<int> > adj(n);
    for (int i = 0; i < n; ++i)
        for (int p = rowptr[i]; p < rowptr[i + 1]; ++p)
        {
            const int j = colidx[p];
            if (j == i || j < 0 || j >= n)
                continue;
            adj[i].push_back(j);
            adj[j].push_back(i);
        }
    const size_t dense_limit = (size_t)std::max(16.0, 10.0 * std::sqrt((double)n));
    std::vector<char> dense(n, 0), done(n, 0);
    for (int i = 0; i < n; ++i)
    {
        std::sort(adj[i].begin(), adj[i].end());
        adj[i].erase(std::unique(adj[i].begin(), adj[i].end()), adj[i].end());
        dense[i] = adj[i].size() > dense_limit;
    }
    std::vector<std::pair<size_t, int> > last;
    for (int i = 0; i < n; ++i)
    {
        if (dense[i])
        {
            last.push_back(std::make_pair(adj[i].size(), i));
            adj[i].clear();
            continue;
        }
        adj[i].erase(std::remove_if(adj[i].begin(), adj[i].end(), [&](int j) { return dense[j] != 0; }),
                     adj[i].end());
    }

    typedef std::pair<size_t, int> Entry;
    std::priority_queue<Entry, std::vector<Entry>, std::greater<Entry> > heap;
    for (int i = 0; i < n; ++i)
        if (!dense[i])
            heap.push(Entry(adj[i].size(), i));

    std::vector<int> order;
    order.reserve(n);
    std::vector<int> merged;
    size_t remaining = 0;
    for (int i = 0; i < n; ++i)
        remaining += !dense[i];
    while (!heap.empty())
    {
        const Entry e = heap.top();
        heap.pop();
        const int v = e.second;
        if (done[v] || e.first != adj[v].size())
            continue;                                  // stale entry
        if (e.first + 1 == remaining)
        {
            // the minimum degree is that of a complete graph: whatever is left
            // is one clique and any order fills it completely
            for (int i = 0; i < n; ++i)
                if (!done[i] && !dense[i])
                {
                    done[i] = 1;
                    order.push_back(i);
                }
            break;
        }
        done[v] = 1;
        --remaining;
        order.push_back(v);
        std::vector<int> nb;
        nb.swap(adj[v]);
        // the neighbours of v become a clique
        for (size_t a = 0; a < nb.size(); ++a)
        {
            const int u = nb[a];
            std::vector<int> &au = adj[u];
            merged.clear();
            size_t i = 0, j = 0;
            while (i < au.size() || j < nb.size())
            {
                int w;
                if (j == nb.size() || (i < au.size() && au[i] < nb[j]))
                    w = au[i++];
                else if (i == au.size() || nb[j] < au[i])
                    w = nb[j++];
                else
                {
                    w = au[i++];
                    ++j;
                }
                if (w != u && w != v)
                    merged.push_back(w);
            }
            au.assign(merged.begin(), merged.end());
            heap.push(Entry(au.size(), u));
        }
    }
    std::sort(last.begin(), last.end());
    for (size_t i = 0; i < last.size(); ++i)
        order.push_back(last[i].second);
    return order;
}

namespace
{

// symmetric adjacency of A + A' without the diagonal; rows with more than
// `dense_limit` neighbours are taken out (flag set, lists cleared)
void symmetric_graph(int n, const int *rowptr, const int *colidx, size_t dense_limit,
                     std::vector<std::vector<int> > &adj, std::vector<char> &dense)
{
    adj.assign(n, std::vector<int>());
    for (int i = 0; i < n; ++i)
        for (int p = rowptr[i]; p < rowptr[i + 1]; ++p)
        {
            const int j = colidx[p];
            if (j == i || j < 0 || j >= n)
                continue;
            adj[i].push_back(j);
            adj[j].push_back(i);
        }
    dense.assign(n, 0);
    for (int i = 0; i < n; ++i)
    {
        std::sort(adj[i].begin(), adj[i].end());
        adj[i].erase(std::unique(adj[i].begin(), adj[i].end()), adj[i].end());
        dense[i] = adj[i].size() > dense_limit;
    }
    for (int i = 0; i < n; ++i)
    {
        if (dense[i])
            continue;
        adj[i].erase(std::remove_if(adj[i].begin(), adj[i].end(), [&](int j) { return dense[j] != 0; }),
                     adj[i].end());
    }
}

}

// Nested dissection by level structures (George): a connected piece is split by
// the middle level of a breadth-first search from a pseudo-peripheral node; the
// two sides are ordered first (recursively), the separator last; pieces of at
// most `leaf` nodes are ordered by minimum degree.  O(n log n) breadth-first
// work, where the explicit-graph minimum degree above is quadratic in the size of
// the last cliques; on grid-like problems the fill is comparable.
std::vector<int> nested_dissection_order(int n, const int *rowptr, const int *colidx, int leaf)
{
    const size_t dense_limit = (size_t)std::max(16.0, 10.0 * std::sqrt((double)n));
    std::vector<std::vector<int> > adj;
    std::vector<char> dense;
    symmetric_graph(n, rowptr, colidx, dense_limit, adj, dense);

    std::vector<int> order;
    order.reserve(n);
    std::vector<int> part(n, -1);          // id of the piece a node currently belongs to
    std::vector<int> dist(n, -1), local(n, -1);
    int next_part = 0;

    // breadth-first search inside piece `id` from `root`; fills `visit` level by level
    auto bfs = [&](int root, int id, std::vector<int> &visit, std::vector<int> &level_start) {
        visit.clear();
        level_start.clear();
        visit.push_back(root);
        dist[root] = 0;
        size_t head = 0;
        while (head < visit.size())
        {
            level_start.push_back((int)head);
            const size_t end = visit.size();
            for (; head < end; ++head)
            {
                const int v = visit[head];
                for (size_t a = 0; a < adj[v].size(); ++a)
                {
                    const int u = adj[v][a];
                    if (part[u] == id && dist[u] < 0)
                    {
                        dist[u] = dist[v] + 1;
                        visit.push_back(u);
                    }
                }
            }
        }
        level_start.push_back((int)visit.size());
        for (size_t i = 0; i < visit.size(); ++i)
            dist[visit[i]] = -1;
    };

    auto order_leaf = [&](std::vector<int> const &nodes) {
        // minimum degree on the induced subgraph
        const int m = (int)nodes.size();
        for (int i = 0; i < m; ++i)
            local[nodes[i]] = i;
        std::vector<int> rp(m + 1, 0), ci;
        for (int i = 0; i < m; ++i)
        {
            const int v = nodes[i];
            for (size_t a = 0; a < adj[v].size(); ++a)
                if (local[adj[v][a]] >= 0)
                    ci.push_back(local[adj[v][a]]);
            rp[i + 1] = (int)ci.size();
        }
        std::vector<int> o = minimum_degree_order(m, rp.data(), ci.data());
        for (int i = 0; i < m; ++i)
            order.push_back(nodes[o[i]]);
        for (int i = 0; i < m; ++i)
            local[nodes[i]] = -1;
    };

    // explicit stack of work items: a piece to dissect, or a separator to emit
    struct Item { std::vector<int> nodes; bool emit; };
    std::vector<Item> stack;
    {
        Item all;
        all.emit = false;
        for (int i = 0; i < n; ++i)
            if (!dense[i])
                all.nodes.push_back(i);
        stack.push_back(std::move(all));
    }
    std::vector<int> visit, level_start, visit2, level_start2;
    while (!stack.empty())
    {
        Item item = std::move(stack.back());
        stack.pop_back();
        if (item.emit)
        {
            order.insert(order.end(), item.nodes.begin(), item.nodes.end());
            continue;
        }
        if (item.nodes.empty())
            continue;
        if ((int)item.nodes.size() <= leaf)
        {
            order_leaf(item.nodes);
            continue;
        }
        const int id = next_part++;
        for (size_t i = 0; i < item.nodes.size(); ++i)
            part[item.nodes[i]] = id;
        // connected component of the first node
        bfs(item.nodes[0], id, visit, level_start);
        if (visit.size() < item.nodes.size())
        {
            // not connected: this component and the rest are independent pieces
            Item comp, rest;
            comp.emit = rest.emit = false;
            for (size_t i = 0; i < visit.size(); ++i)
                local[visit[i]] = 1;
            for (size_t i = 0; i < item.nodes.size(); ++i)
                (local[item.nodes[i]] == 1 ? comp : rest).nodes.push_back(item.nodes[i]);
            for (size_t i = 0; i < visit.size(); ++i)
                local[visit[i]] = -1;
            stack.push_back(std::move(rest));
            stack.push_back(std::move(comp));
            continue;
        }
        // pseudo-peripheral node: restart from a node of the last level while the depth grows
        for (int tries = 0; tries < 4; ++tries)
        {
            const int last = visit.back();
            bfs(last, id, visit2, level_start2);
            const bool deeper = level_start2.size() > level_start.size();
            visit.swap(visit2);
            level_start.swap(level_start2);
            if (!deeper)
                break;
        }
        const int depth = (int)level_start.size() - 1;      // number of levels
        if (depth < 3)
        {
            order_leaf(item.nodes);                        // a clique-like piece: nothing to cut
            continue;
        }
        // the level where the running count crosses one half, never the first or last one
        int cut = 1;
        while (cut < depth - 2 && level_start[cut + 1] < (int)visit.size() / 2)
            ++cut;
        Item left, right, sep;
        left.emit = right.emit = false;
        sep.emit = true;
        left.nodes.assign(visit.begin(), visit.begin() + level_start[cut]);
        sep.nodes.assign(visit.begin() + level_start[cut], visit.begin() + level_start[cut + 1]);
        right.nodes.assign(visit.begin() + level_start[cut + 1], visit.end());
        stack.push_back(std::move(sep));                   // popped last: ordered after both sides
        stack.push_back(std::move(right));
        stack.push_back(std::move(left));
    }
    // dense rows last, lightest first
    std::vector<std::pair<size_t, int> > last;
    for (int i = 0; i < n; ++i)
        if (dense[i])
        {
            size_t deg = 0;
            for (int p = rowptr[i]; p < rowptr[i + 1]; ++p)
                ++deg;
            last.push_back(std::make_pair(deg, i));
        }
    std::sort(last.begin(), last.end());
    for (size_t i = 0; i < last.size(); ++i)
        order.push_back(last[i].second);
    return order;
}

// Entries of the Cholesky factor of the pattern of A + A' under `order` (symbolic
// factorisation by merging the structures of the children in the elimination
// tree): the fill an LU with diagonal pivots would have in L and in U.
long long symbolic_fill(int n, const int *rowptr, const int *colidx, const int *order)
{
    std::vector<std::vector<int> > adj;
    std::vector<char> dense;
    symmetric_graph(n, rowptr, colidx, (size_t)n + 1, adj, dense);
    std::vector<int> pos(n);
    for (int j = 0; j < n; ++j)
        pos[order[j]] = j;
    std::vector<std::vector<int> > structure(n), children(n);
    std::vector<int> merged, tmp;
    long long fill = 0;
    for (int j = 0; j < n; ++j)
    {
        const int v = order[j];
        merged.clear();
        for (size_t a = 0; a < adj[v].size(); ++a)
            if (pos[adj[v][a]] > j)
                merged.push_back(pos[adj[v][a]]);
        std::sort(merged.begin(), merged.end());
        for (size_t c = 0; c < children[j].size(); ++c)
        {
            std::vector<int> &sc = structure[children[j][c]];
            tmp.clear();
            size_t x = 0, y = 0;
            while (x < merged.size() || y < sc.size())
            {
                int w;
                if (y == sc.size() || (x < merged.size() && merged[x] < sc[y]))
                    w = merged[x++];
                else if (x == merged.size() || sc[y] < merged[x])
                    w = sc[y++];
                else
                {
                    w = merged[x++];
                    ++y;
                }
                if (w > j)
                    tmp.push_back(w);
            }
            merged.swap(tmp);
            std::vector<int>().swap(sc);           // the child's structure is not needed again
        }
        fill += (long long)merged.size();
        if (!merged.empty())
            children[merged[0]].push_back(j);
        structure[j] = merged;
    }
    return fill;
}

int SparseLU::factor(int n_in, const int *rowptr, const int *colidx, const double *vals,
                     double pivot_threshold, const int *order)
{
    n = n_in;
    const int nnz_a = n > 0 ? rowptr[n] : 0;

    // column access to A
    std::vector<int> Ap(n + 1, 0), Ai(nnz_a);
    std::vector<double> Ax(nnz_a);
    for (int p = 0; p < nnz_a; ++p)
        ++Ap[colidx[p] + 1];
    for (int j = 0; j < n; ++j)
        Ap[j + 1] += Ap[j];
    {
        std::vector<int> next(Ap.begin(), Ap.end() - 1);
        for (int i = 0; i < n; ++i)
            for (int p = rowptr[i]; p < rowptr[i + 1]; ++p)
            {
                const int dst = next[colidx[p]]++;
                Ai[dst] = i;
                Ax[dst] = vals[p];
            }
    }

    if (order)
        q.assign(order, order + n);
    else
    {
        // nested dissection wins on mesh-like graphs (less fill, bushier elimination
        // tree, O(n log n)), minimum degree on irregular ones; where the latter is
        // affordable both are computed and the one with the smaller symbolic fill taken
        q = nested_dissection_order(n, rowptr, colidx);
        if (n <= compare_orderings_up_to)
        {
            std::vector<int> md = minimum_degree_order(n, rowptr, colidx);
            if (symbolic_fill(n, rowptr, colidx, md.data()) < symbolic_fill(n, rowptr, colidx, q.data()))
                q.swap(md);
        }
    }

    pinv.assign(n, -1);
    Lp.assign(1, 0);
    Up.assign(1, 0);
    Li.clear(); Lx.clear(); Ui.clear(); Ux.clear();
    off_diagonal_pivots = 0;
    min_abs_pivot = 0.0;
    max_abs_pivot = 0.0;

    std::vector<double> x(n, 0.0);
    std::vector<int> mark(n, -1), topo(n), stack(n), resume(n);

    for (int j = 0; j < n; ++j)
    {
        const int c = q[j];

        // rows reached from the pattern of A(:, c) through the columns of L
        // found so far, in topological order (topo[top .. n))
        int top = n;
        for (int pa = Ap[c]; pa < Ap[c + 1]; ++pa)
        {
            if (mark[Ai[pa]] == j)
                continue;
            int head = 0;
            stack[0] = Ai[pa];
            while (head >= 0)
            {
                const int i = stack[head];
                const int col = pinv[i];
                if (mark[i] != j)
                {
                    mark[i] = j;
                    resume[head] = col < 0 ? 0 : Lp[col];
                }
                bool finished = true;
                if (col >= 0)
                {
                    const int pend = Lp[col + 1];
                    for (int p = resume[head]; p < pend; ++p)
                    {
                        const int r = Li[p];
                        if (mark[r] == j)
                            continue;
                        resume[head] = p + 1;
                        stack[++head] = r;
                        finished = false;
                        break;
                    }
                }
                if (finished)
                {
                    --head;
                    topo[--top] = i;
                }
            }
        }

        for (int pa = Ap[c]; pa < Ap[c + 1]; ++pa)
            x[Ai[pa]] += Ax[pa];

        // x := L^-1 x on that pattern; the pivotal rows give U(:, j)
        for (int t = top; t < n; ++t)
        {
            const int i = topo[t];
            const int col = pinv[i];
            if (col < 0)
                continue;
            const double xi = x[i];
            Ui.push_back(col);
            Ux.push_back(xi);
            if (xi != 0.0)
                for (int p = Lp[col]; p < Lp[col + 1]; ++p)
                    x[Li[p]] -= Lx[p] * xi;
        }

        // threshold partial pivoting with preference for the diagonal
        double biggest = 0.0;
        int pivot_row = -1;
        for (int t = top; t < n; ++t)
        {
            const int i = topo[t];
            if (pinv[i] < 0 && std::fabs(x[i]) > biggest)
            {
                biggest = std::fabs(x[i]);
                pivot_row = i;
            }
        }
        if (pivot_row < 0 || biggest == 0.0)
        {
            for (int t = top; t < n; ++t)
                x[topo[t]] = 0.0;
            return j + 1;
        }
        if (pinv[c] < 0 && mark[c] == j && std::fabs(x[c]) >= pivot_threshold * biggest && x[c] != 0.0)
            pivot_row = c;
        else
            ++off_diagonal_pivots;
        const double pivot = x[pivot_row];
        const double ap = std::fabs(pivot);
        if (j == 0 || ap < min_abs_pivot) min_abs_pivot = ap;
        if (j == 0 || ap > max_abs_pivot) max_abs_pivot = ap;

        Ui.push_back(j);                 // the diagonal is the last entry of its column
        Ux.push_back(pivot);
        Up.push_back((int)Ui.size());

        pinv[pivot_row] = j;
        for (int t = top; t < n; ++t)
        {
            const int i = topo[t];
            if (pinv[i] < 0 && x[i] != 0.0)
            {
                Li.push_back(i);         // original row id until the end
                Lx.push_back(x[i] / pivot);
            }
            x[i] = 0.0;
        }
        Lp.push_back((int)Li.size());
    }

    for (size_t p = 0; p < Li.size(); ++p)
        Li[p] = pinv[Li[p]];
    return 0;
}

void SparseLU::solve(double *xs, int nrhs, long long ld, bool trans) const
{
    std::vector<double> y(n);
    for (int r = 0; r < nrhs; ++r)
    {
        double *x = xs + (long long)r * ld;
        if (!trans)
        {
            for (int i = 0; i < n; ++i)
                y[pinv[i]] = x[i];
            for (int j = 0; j < n; ++j)
            {
                const double yj = y[j];
                for (int p = Lp[j]; p < Lp[j + 1]; ++p)
                    y[Li[p]] -= Lx[p] * yj;
            }
            for (int j = n - 1; j >= 0; --j)
            {
                const int pd = Up[j + 1] - 1;
                y[j] /= Ux[pd];
                const double yj = y[j];
                for (int p = Up[j]; p < pd; ++p)
                    y[Ui[p]] -= Ux[p] * yj;
            }
            for (int j = 0; j < n; ++j)
                x[q[j]] = y[j];
        }
        else
        {
            for (int j = 0; j < n; ++j)
                y[j] = x[q[j]];
            for (int j = 0; j < n; ++j)
            {
                const int pd = Up[j + 1] - 1;
                double s = y[j];
                for (int p = Up[j]; p < pd; ++p)
                    s -= Ux[p] * y[Ui[p]];
                y[j] = s / Ux[pd];
            }
            for (int j = n - 1; j >= 0; --j)
            {
                double s = y[j];
                for (int p = Lp[j]; p < Lp[j + 1]; ++p)
                    s -= Lx[p] * y[Li[p]];
                y[j] = s;
            }
            for (int i = 0; i < n; ++i)
                x[i] = y[pinv[i]];
        }
    }
}

}
