;; UNVERIFIED: no JVM on the box this was written on; it has never been compiled.
;; `infer :smc` on the GPU population engine.  Replaces src/anglican/smc.cljc:53-63 (the
;; `infer` method) and :35-51 (`sweep`): one call of apop_smc_sweep is one sweep - N particles
;; run to every global observe, systematic resampling at each (smc.cljc:41-43, 86-137).
;; The returned states have the shape inference.cljc:114-117 (`stripdown`) leaves: every state
;; of a sweep carries the same log-weight, the sweep's log marginal likelihood estimate.
(ns anglican.gpu.smc
  (:refer-clojure :exclude [rand rand-int rand-nth])
  (:require [anglican.inference :refer [infer]]
            [anglican.state :as st]
            [anglican.gpu.ffi :as ffi])
  (:import [java.lang.foreign Arena MemorySegment ValueLayout]))

(defn create
  "one population of n particles on `device`; seed keys the counter-based generator that
  replaces the reference's global synchronised one (runtime.cljc:142-147)"
  ([n] (create n 0 (System/nanoTime)))
  ([n device seed]
   (with-open [arena (Arena/ofConfined)]
     (let [out (.allocate arena ValueLayout/ADDRESS)
           rc (.invokeWithArguments ffi/apop-create
                                    [out (long n) (long seed) (int device) (int 0) (int 1)
                                     MemorySegment/NULL])]
       (when (neg? rc)
         (throw (ex-info "apop_create failed (no CUDA device? there is no CPU fallback)" {:code rc})))
       (.get out ValueLayout/ADDRESS 0)))))

(defn- doubles->segment [^Arena arena xs]
  (.allocateFrom arena ValueLayout/JAVA_DOUBLE (double-array xs)))

(defn ssm-segment
  "struct apop_ssm for a state-space program.  `model` is
  {:kind :lgssm1d|:sv|:hmm|:lgmvn, ...parameters named as the C fields}"
  [^Arena arena model]
  (let [seg (.allocate arena ffi/apop-ssm-layout)
        kinds {:lgssm1d 1 :sv 2 :hmm 3 :lgmvn 4}]
    (.set seg ValueLayout/JAVA_INT 0 (int (kinds (:kind model))))
    (.set seg ValueLayout/JAVA_INT 4 (int (:dim model 1)))
    (doseq [[i k] (map-indexed vector [:m0 :s0 :a :b :q :c :d :r :beta])]
      (.set seg ValueLayout/JAVA_DOUBLE (long (+ 8 (* 8 i))) (double (get model k 0.0))))
    (doseq [[i k] (map-indexed vector [:hmm-init :hmm-trans :hmm-mu :hmm-sd :A :Lq :R])]
      (.set seg ValueLayout/ADDRESS (long (+ 80 (* 8 i)))
            (if-let [v (get model k)] (doubles->segment arena (flatten v)) MemorySegment/NULL)))
    seg))

(defn sweep
  "one `sweep` (smc.cljc:35-51).  opts: the fields of apop_sweep_opts as keywords
  (:csmc :n-cond-steps :reset-weights :keep-history :no-resample :ancestor-sampling ...).
  Returns {:log-z double :states double[n*dim]}"
  [ctx n model ys & {:keys [retained] :as opts}]
  (with-open [arena (Arena/ofConfined)]
    (let [dim (:dim model 1)
          y (doubles->segment arena ys)
          o (.allocate arena ffi/apop-sweep-opts-layout)
          r (.allocate arena ffi/apop-sweep-result-layout)
          xs (.allocate arena ValueLayout/JAVA_DOUBLE (long (* n dim)))
          field {:csmc 0 :n-cond-steps 4 :reset-weights 8 :keep-history 12 :use-graph 16
                 :want-filt 24 :no-resample 28 :ancestor-sampling 32}]
      (assert (= (.byteSize ffi/apop-sweep-opts-layout)
                 (.invokeWithArguments ffi/apop-sizeof-sweep-opts []))
              "apop_sweep_opts layout differs from the loaded library")
      (.fill o (byte 0))
      (.fill r (byte 0))
      (doseq [[k off] field] (.set o ValueLayout/JAVA_INT (long off) (int (get opts k 0))))
      (.set o ValueLayout/JAVA_INT 20 (int 1))                       ; want_final
      (.set r ValueLayout/ADDRESS 32 xs)                             ; x_final
      (ffi/call ffi/apop-smc-sweep ctx (ssm-segment arena model) y
                (int (quot (count ys) dim))
                (if retained (doubles->segment arena retained) MemorySegment/NULL) o r)
      {:log-z (.get r ValueLayout/JAVA_DOUBLE 0)
       :states (.toArray xs ValueLayout/JAVA_DOUBLE)})))

(defn states->seq
  "the `(map :state particles)` of smc.cljc:61 for a state-space program: one state per
  particle, log-weight = the sweep's log Z, :result = the particle's final latent state"
  [{:keys [log-z states]} dim]
  (map (fn [x] (-> st/initial-state
                   (st/set-log-weight log-z)
                   (assoc :result (if (= dim 1) (first x) (vec x)))))
       (partition dim states)))

(defmethod infer :smc [_ prog value
                       & {:keys [number-of-particles device] ; per sweep
                          :or {number-of-particles 1 device 0}}]
  (assert (>= number-of-particles 1)
          ":number-of-particles must be at least 1")                 ; smc.cljc:56-57
  ;; `prog` is a state-space program description {:model ... :observations ...}; a general
  ;; lock-step program is traced once per step over anglican.gpu.runtime's batched records
  ;; and resampled with runtime/resample! at every global observe - a traced control flow
  ;; that would depend on one particle's value is where the reference's
  ;; "some `observe' directives are not global" (smc.cljc:48-51) is raised.
  (let [{:keys [model observations]} (if (fn? prog) (prog value) prog)
        ctx (create number-of-particles device (System/nanoTime))
        sweep-no (atom 0)]
    (letfn [(sample-seq []
              (lazy-seq
                (.invokeWithArguments ffi/apop-set-seed [ctx (long (swap! sweep-no inc))])
                (let [out (sweep ctx number-of-particles model observations)]
                  (concat (states->seq out (:dim model 1)) (sample-seq)))))]
      (sample-seq))))
