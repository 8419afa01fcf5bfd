;; UNVERIFIED: no JVM on the box this was written on; it has never been compiled.
;; `infer :ipmcmc` - interacting particle MCMC.  Replaces src/anglican/ipmcmc.cljc:148-208:
;; M nodes, P of them conditional; each node is one population (one apop_ctx, several per GPU)
;; swept from the claypoole pool exactly as the reference does (ipmcmc.cljc:176-182); the only
;; data that crosses nodes is the vector of node log-marginals (ipmcmc.cljc:183-187), on which
;; apop_ipmcmc_gibbs redraws the conditional indices (`gibbs-update-csmc-indices`,
;; ipmcmc.cljc:106-146, with the intended sampling rule - see DESIGN.md 3.5).
(ns anglican.gpu.ipmcmc
  (:require [anglican.inference :refer [infer]]
            [anglican.runtime :refer [sample* uniform-continuous]]
            [com.climate.claypoole :as cp]
            [anglican.gpu.ffi :as ffi]
            [anglican.gpu.smc :as smc]
            [anglican.gpu.pgibbs :as pg])
  (:import [java.lang.foreign Arena ValueLayout]))

(defn gibbs-update
  "log-Zs of all M nodes, P conditional -> {:csmc-idx [P node indices] :zeta-sums [M]}"
  [log-zs p]
  (with-open [arena (Arena/ofConfined)]
    (let [m (count log-zs)
          lz (.allocateFrom arena ValueLayout/JAVA_DOUBLE (double-array log-zs))
          us (.allocateFrom arena ValueLayout/JAVA_DOUBLE
                            (double-array (repeatedly p #(sample* (uniform-continuous 0. 1.)))))
          idx (.allocate arena ValueLayout/JAVA_INT (long p))
          zs (.allocate arena ValueLayout/JAVA_DOUBLE (long m))
          rc (.invokeWithArguments ffi/apop-ipmcmc-gibbs [lz (int m) (int p) us idx zs])]
      (assert (zero? rc))
      {:csmc-idx (vec (.toArray idx ValueLayout/JAVA_INT))
       :zeta-sums (vec (.toArray zs ValueLayout/JAVA_DOUBLE))})))

(defmethod infer :ipmcmc [_ prog value
                          & {:keys [number-of-particles number-of-nodes number-of-csmc-nodes
                                    pool devices all-particles?]
                             :or {number-of-particles 2 number-of-nodes 32 pool :builtin
                                  devices [0] all-particles? true}}]
  (assert (>= number-of-particles 2)
          ":number-of-particles must be at least 2")                 ; ipmcmc.cljc:160-161
  (let [number-of-csmc-nodes (or number-of-csmc-nodes (quot number-of-nodes 2))
        {:keys [model observations]} (if (fn? prog) (prog value) prog)
        dim (:dim model 1)
        T (quot (count observations) dim)
        ;; node i lives on devices[i mod #devices]: 4 nodes per GPU in BASELINE config 4
        ctxs (mapv #(smc/create number-of-particles (nth devices (mod % (count devices)))
                                (+ (System/nanoTime) %))
                   (range number-of-nodes))
        node-sweep (fn [i retained iter]
                     (let [ctx (ctxs i)]
                       (.invokeWithArguments ffi/apop-set-seed
                                             [ctx (long (+ (* iter number-of-nodes) i))])
                       (smc/sweep ctx number-of-particles model observations
                                  :retained retained :csmc (if retained 1 0)
                                  :n-cond-steps (dec T) :reset-weights 1 :keep-history 2)))]
    (letfn [(sample-seq [retained-paths iter]               ; node index -> retained path or nil
              (lazy-seq
                (let [outs (vec (cp/pmap pool #(node-sweep % (get retained-paths %) iter)
                                         (range number-of-nodes)))
                      {:keys [csmc-idx]} (gibbs-update (map :log-z outs) number-of-csmc-nodes)
                      ;; each conditional slot continues from one lineage of the node it drew
                      retained' (into {}
                                      (map-indexed
                                        (fn [j node]
                                          [j (pg/trace-lineage (ctxs node)
                                                               (rand-int number-of-particles) T dim)])
                                        csmc-idx))
                      emit (if all-particles? outs (map outs csmc-idx))]
                  (concat (mapcat #(smc/states->seq % dim) emit)
                          (sample-seq retained' (inc iter))))))]
      (sample-seq {} 0))))
