-- BASELINE config C1 through the reference's own aggregate (shape of
-- examples/performance_10k_groups_R/performance_test_ols_aggregate.sql:53-66): ols_fit_agg GROUP BY over
-- 10 000 groups x 100 rows x 3 features, f64, synthetic.
-- The reference's Finalize writes every result list with ListVector::SetListSize and never reserves the child vector
-- (SetListInResult, src/aggregate_functions/ols_aggregate.cpp:237-246): a result chunk of more than 2048 / p groups
-- overruns the child's 2048-entry buffer and corrupts the heap (glibc aborts the shell). The sources are used
-- unmodified, so the query is issued in slices of 600 groups (1 800 coefficients per chunk at most); the times add.
PRAGMA threads=__THREADS__;
CREATE TABLE t AS
SELECT (i % 10000)::INTEGER AS group_id,
       (random() * 20 - 10) AS x1, (random() * 20 - 10) AS x2, (random() * 20 - 10) AS x3,
       0.0::DOUBLE AS y
FROM range(1000000) r(i);
UPDATE t SET y = 1.5 + 2.0 * x1 - 3.0 * x2 + 0.5 * x3 + (random() - 0.5) * 4.0;
CREATE TABLE r1 (group_id INTEGER, r2 DOUBLE, n BIGINT, b1 DOUBLE);
CREATE TABLE r3 (group_id INTEGER, r2 DOUBLE, p1 DOUBLE);
.timer on
__SLICES__
.timer off
SELECT count(*) AS groups, avg(r2) AS mean_r2, min(n) AS min_n, avg(b1) AS mean_b1 FROM r1;
SELECT count(*) AS groups, max(p1) AS max_p1 FROM r3;
