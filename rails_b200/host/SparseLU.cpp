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
        q = minimum_degree_order(n, rowptr, colidx);

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
