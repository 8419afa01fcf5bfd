;; UNVERIFIED: no JVM on the box this was written on; it has never been compiled.
;; Batched `distribution` records: the same protocol as anglican.runtime
;; (src/anglican/runtime.cljc:119-124, `sample*` / `observe*`), but a value is a COLUMN of
;; one number per particle living in GPU memory, and `observe*` adds into the population's
;; log-weight column instead of returning a double (smc.cljc:23-27, state.cljc:27-31).
(ns anglican.gpu.runtime
  (:require [anglican.runtime :refer [distribution]]
            [anglican.gpu.ffi :as ffi])
  (:import [java.lang.foreign Arena MemorySegment ValueLayout]))

(defrecord Column [ctx handle dtype width])

(defn column-alloc
  "a device column of `width` values per particle; dtype :f64 or :i32"
  [ctx dtype width]
  (with-open [arena (Arena/ofConfined)]
    (let [out (.allocate arena ValueLayout/JAVA_INT)]
      (ffi/call ffi/apop-column-alloc ctx (int (if (= dtype :i32) 1 0)) (int width) out)
      (->Column ctx (.get out ValueLayout/JAVA_INT 0) dtype width))))

(defn arg
  "number or Column -> struct apop_arg {col, scalar}; col < 0 means `scalar` for every particle"
  [^Arena arena v]
  (let [seg (.allocate arena ffi/apop-arg-layout)]
    (if (instance? Column v)
      (.set seg ValueLayout/JAVA_INT 0 (int (:handle v)))
      (do (.set seg ValueLayout/JAVA_INT 0 (int -1))
          (.set seg ValueLayout/JAVA_DOUBLE 8 (double v))))
    seg))

(defmacro defbatched
  "a two-parameter scalar family: `sample*` fills and returns a fresh column, `observe*`
  accumulates log p(value) into `logw` and returns that column"
  [record-name sample-handle observe-handle]
  `(defrecord ~record-name [~'ctx ~'p1 ~'p2 ~'draw-id ~'logw]
     distribution
     (~'sample* [~'_]
       (with-open [arena# (Arena/ofConfined)]
         (let [out# (column-alloc ~'ctx :f64 1)]
           (ffi/call ~sample-handle ~'ctx (arg arena# ~'p1) (arg arena# ~'p2) (int ~'draw-id)
                     (int (:handle out#)))
           out#)))
     (~'observe* [~'_ value#]
       (with-open [arena# (Arena/ofConfined)]
         (ffi/call ~observe-handle ~'ctx (arg arena# ~'p1) (arg arena# ~'p2) (arg arena# value#)
                   (int (:handle ~'logw)))
         ~'logw))))

(defbatched batched-normal ffi/apop-sample-normal ffi/apop-observe-normal)            ; runtime.cljc:403-409
(defbatched batched-gamma ffi/apop-sample-gamma ffi/apop-observe-gamma)               ; :325-330 (shape, rate)
(defbatched batched-beta ffi/apop-sample-beta ffi/apop-observe-beta)                  ; :251-256
(defbatched batched-uniform-continuous ffi/apop-sample-uniform ffi/apop-observe-uniform) ; :450-453

;; discrete / dirichlet / mvn / mvt / wishart take host parameter arrays (weights, alpha, mean +
;; covariance): same shape of record around apop-sample-discrete ... apop-observe-wishart; the
;; nine remaining scalar families go through apop-sample-dist / apop-observe-dist with the
;; family code of include/anglican_pop.h (APOP_DIST_*).

(defn normalize
  "particle-weights / log-marginal / ESS of a log-weight column (smc.cljc:70-84,
  inference.cljc:181-186): {:max :log-mean-w :ess}"
  [ctx logw]
  (with-open [arena (Arena/ofConfined)]
    (let [m (.allocate arena ValueLayout/JAVA_DOUBLE)
          l (.allocate arena ValueLayout/JAVA_DOUBLE)
          e (.allocate arena ValueLayout/JAVA_DOUBLE)]
      (ffi/call ffi/apop-normalize ctx (int (:handle logw)) m l e)
      {:max (.get m ValueLayout/JAVA_DOUBLE 0)
       :log-mean-w (.get l ValueLayout/JAVA_DOUBLE 0)
       :ess (.get e ValueLayout/JAVA_DOUBLE 0)})))

(defn resample!
  "`resample` of smc.cljc:86-137 on the device: ancestors by the systematic wheel, then every
  listed column is replaced by its offspring copy (smc.cljc:124-129).  Returns the common
  log-weight of the offspring (log(step) + max, or -inf: smc.cljc:98-103)."
  [ctx logw anc-col n-new cols]
  (with-open [arena (Arena/ofConfined)]
    (let [new-lw (.allocate arena ValueLayout/JAVA_DOUBLE)
          hs (.allocateFrom arena ValueLayout/JAVA_INT (int-array (map :handle cols)))]
      (ffi/call ffi/apop-resample ctx (int (:handle logw)) (int 0) Double/NaN (long n-new)
                (int 1) (int (:handle anc-col)) new-lw MemorySegment/NULL)
      (ffi/call ffi/apop-gather ctx (int (:handle anc-col)) (long n-new) hs (int (count cols)))
      (.get new-lw ValueLayout/JAVA_DOUBLE 0))))
